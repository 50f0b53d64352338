// Micro-benchmark (development aid): DFMA rate of a 4x8 register tile whose three operands are all vector registers
// (acc[r][j] = fma(w[j], a[r], acc[r][j]), the inner product step of the tiled kernels), against the same number of
// DFMAs whose multiplier is a uniform register (the specialised kernels).  No memory traffic in the loop.
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o ubench_tile ubench_tile.cu
#include <cstdio>
#include <cuda_runtime.h>
struct W { double w[64]; };
template <int MODE>
__global__ void __launch_bounds__(256) k(const __grid_constant__ W p, int iters, double *sink, const double *src)
{
    double acc[4][8], a[4], w[8];
#pragma unroll
    for (int r = 0; r < 4; ++r) { a[r] = src[threadIdx.x + r];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[r][j] = r + j; }
#pragma unroll
    for (int j = 0; j < 8; ++j) w[j] = src[threadIdx.x + 8 + j];
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    acc[r][j] = __fma_rn(MODE == 0 ? w[j] : p.w[(u * 8 + j) & 63], a[r], acc[r][j]);
            // rotate the operands so that nothing is loop-invariant per accumulator
            const double t = a[0]; a[0] = a[1]; a[1] = a[2]; a[2] = a[3]; a[3] = t;
        }
    }
    double s = 0;
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int j = 0; j < 8; ++j) s += acc[r][j];
    if (s == 12345.678) sink[0] = s;
}
template <int MODE> void run(const char *name, int warps_per_smsp)
{
    W h; for (int i = 0; i < 64; ++i) h.w[i] = 1.0 + 1e-9 * i;
    double *sink, *src; cudaMalloc(&sink, 8); cudaMalloc(&src, 8 * 1024); cudaMemset(src, 0, 8 * 1024);
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 4000, threads = 128 * warps_per_smsp > 256 ? 256 : 128 * warps_per_smsp;
    const int grid = sms * (128 * warps_per_smsp / threads);
    k<MODE><<<grid, threads>>>(h, 100, sink, src);
    cudaEventRecord(e0); k<MODE><<<grid, threads>>>(h, iters, sink, src); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double tf = (double)grid * threads * iters * 4 * 32 * 2 / (ms * 1e-3) / 1e12;
    printf("%-4s warps/scheduler %d: %.2f TFLOP/s (%s)\n", name, warps_per_smsp, tf, cudaGetErrorString(cudaGetLastError()));
}
int main()
{
    for (int w : {1, 2, 4}) { run<0>("RRR", w); run<1>("RUR", w); }
    return 0;
}
