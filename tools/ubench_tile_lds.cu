// Micro-benchmark (development aid): the k-loop of fb_profile_tile_kernel in isolation -- 4x8 register tile, operands
// from shared memory (activations [k][TM], weights [k][OUTP]), 8 warps per CTA, one CTA per SM, no barriers, no staging.
// MODE 0: plain loop (unroll 4); MODE 1: ping-pong register prefetch as in the kernel; MODE 2: weights only from smem;
// MODE 3: activations only from smem.
#include <cstdio>
#include <cuda_runtime.h>
constexpr int TM = 64, OUTP = 128, KC = 64;
template <int MODE>
__global__ void __launch_bounds__(256) k(int iters, double *sink)
{
    extern __shared__ __align__(16) double sm[];
    double *A = sm, *Wt = sm + KC * TM;
    for (int i = threadIdx.x; i < KC * TM; i += 256) A[i] = 1e-3 * i;
    for (int i = threadIdx.x; i < KC * OUTP; i += 256) Wt[i] = 1.0 + 1e-6 * i;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, lane_m = lane & 7, lane_n = lane >> 3;
    const int mb = warp & 1, nb = warp >> 1;
    const int m0 = mb * 32 + lane_m * 2, n0 = nb * 32 + lane_n * 8;
    double acc[4][8];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[r][j] = 0.0;
    double2 xa[2] = {}, wa[4] = {};
    for (int it = 0; it < iters; ++it) {
        const double *ap = A + m0, *wp = Wt + n0;
#pragma unroll 4
        for (int kk = 0; kk < KC; ++kk, ap += TM, wp += OUTP) {
            if (MODE != 2) { xa[0] = *reinterpret_cast<const double2 *>(ap); xa[1] = *reinterpret_cast<const double2 *>(ap + 16); }
            if (MODE != 3) {
#pragma unroll
                for (int q = 0; q < 4; ++q) wa[q] = reinterpret_cast<const double2 *>(wp)[q];
            }
            const double a_[4] = {xa[0].x, xa[0].y, xa[1].x, xa[1].y};
            const double w_[8] = {wa[0].x, wa[0].y, wa[1].x, wa[1].y, wa[2].x, wa[2].y, wa[3].x, wa[3].y};
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[r][j] = __fma_rn(w_[j], a_[r], acc[r][j]);
        }
    }
    double s = 0;
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int j = 0; j < 8; ++j) s += acc[r][j];
    if (s == 12345.678) sink[0] = s;
}
template <int MODE> void run(const char *name)
{
    double *sink; cudaMalloc(&sink, 8);
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const size_t smem = sizeof(double) * (KC * TM + KC * OUTP);
    cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 400;
    k<MODE><<<sms, 256, smem>>>(20, sink);
    cudaEventRecord(e0); k<MODE><<<sms, 256, smem>>>(iters, sink); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double tf = (double)sms * 256 * iters * KC * 32 * 2 / (ms * 1e-3) / 1e12;
    printf("%-28s %.2f TFLOP/s (%s)\n", name, tf, cudaGetErrorString(cudaGetLastError()));
}
int main()
{
    run<0>("both operands from smem"); run<2>("weights only from smem"); run<3>("activations only from smem");
    return 0;
}
