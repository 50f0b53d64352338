// Micro-benchmark (development aid): how many DFMAs per uniform constant load (C) and how many resident warps
// per SM sub-partition (W) it takes to keep the FP64 pipe busy when every multiplier comes from the constant bank.
#include <cstdio>
#include <cuda_runtime.h>
struct Wt { double w[512]; };
template <int C>
__global__ void __launch_bounds__(128) k(const __grid_constant__ Wt p, int iters, double *sink)
{
    double a[C], x[C];
#pragma unroll
    for (int c = 0; c < C; ++c) { a[c] = threadIdx.x + c; x[c] = 1e-9 * (threadIdx.x + 3 * c); }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 256; ++u) {
            const double w = p.w[(u * 7) & 511];
#pragma unroll
            for (int c = 0; c < C; ++c) a[c] = __fma_rn(w, x[c], a[c]);
        }
    }
    double s = 0;
#pragma unroll
    for (int c = 0; c < C; ++c) s += a[c];
    if (s == 12345.678) sink[0] = s;
}
template <int C> void run(int warps_per_smsp)
{
    Wt h; for (int i = 0; i < 512; ++i) h.w[i] = 1.0 + 1e-9 * i;
    double *sink; cudaMalloc(&sink, 8);
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 400, grid = sms * warps_per_smsp;      // 128-thread CTAs: one warp per sub-partition each
    k<C><<<grid, 128>>>(h, 20, sink);
    cudaEventRecord(e0); k<C><<<grid, 128>>>(h, iters, sink); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double tf = (double)grid * 128 * iters * 256 * C * 2 / (ms * 1e-3) / 1e12;
    printf("C=%d W=%2d  %.2f TFLOP/s\n", C, warps_per_smsp, tf);
}
int main()
{
    for (int w : {1, 2, 3, 4, 6, 8, 16}) { run<1>(w); run<2>(w); run<4>(w); run<8>(w); }
    return 0;
}
