// Micro-benchmark (development aid): DFMA throughput by operand source on sm_100a.
//   R   : multiplier in a vector register
//   UR  : multiplier fetched from the kernel-parameter constant bank (LDCU -> uniform register operand)
//   RMOV: multiplier fetched from the constant bank, then forced into a vector register
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o ubench_dfma ubench_dfma.cu
#include <cstdio>
#include <cuda_runtime.h>
struct W { double w[256]; };

template <int MODE>
__global__ void __launch_bounds__(256) k(const __grid_constant__ W p, int iters, double *sink)
{
    double a[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) a[c] = threadIdx.x + c;
    double wr = p.w[threadIdx.x & 255];     // per-thread value: must live in a vector register
    double x[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) x[c] = 1e-9 * (threadIdx.x + 3 * c);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 32; ++u) {
            double w;
            if (MODE == 0) w = wr;
            else if (MODE == 1) w = p.w[(u * 7) & 255];
            else { w = p.w[(u * 7) & 255]; asm volatile("mov.f64 %0, %0;" : "+d"(w)); }
#pragma unroll
            for (int c = 0; c < 8; ++c) a[c] = __fma_rn(w, x[c], a[c]);
        }
    }
    double s = 0;
#pragma unroll
    for (int c = 0; c < 8; ++c) s += a[c];
    if (s == 12345.678) sink[0] = s;
}

template <int MODE> void run(const char *name)
{
    W h; for (int i = 0; i < 256; ++i) h.w[i] = 1.0 + 1e-9 * i;
    double *sink; cudaMalloc(&sink, 8);
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 2000, grid = sms * 8;
    k<MODE><<<grid, 256>>>(h, 100, sink);
    cudaEventRecord(e0); k<MODE><<<grid, 256>>>(h, iters, sink); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double tf = (double)grid * 256 * iters * 32 * 8 * 2 / (ms * 1e-3) / 1e12;
    printf("%-5s %.2f ms  %.2f TFLOP/s  (%s)\n", name, ms, tf, cudaGetErrorString(cudaGetLastError()));
}
int main() { run<0>("R"); run<1>("UR"); run<2>("RMOV"); return 0; }
