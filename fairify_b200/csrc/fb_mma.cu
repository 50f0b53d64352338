// fb_mma.cu -- launcher of the DMMA activation-profile kernel (fb_mma.cuh); its own compile unit.
#include <cstdio>
#include <cstdlib>

#include "fb_launch.cuh"
#include "fb_mma.cuh"

// FAIRIFY_B200_MMA_MS = 1 | 2, FAIRIFY_B200_MMA_WARPS = 1..8 and FAIRIFY_B200_MMA_NONRESIDENT=1 override the choices
// below (tuning aids)
static int fb_env_int(const char *name)
{
    const char *v = getenv(name);
    return v ? atoi(v) : 0;
}

template <int MS, int CAP, bool PHILOX, bool RESIDENT>
static cudaError_t fb_mma_launch(const FbMmaNet &net, int warps, size_t smem, int64_t n, const int32_t *boxes,
                                 const int32_t *samples, uint64_t seed, int64_t first_index, int nsamp, int chunks,
                                 int chunk_len, uint32_t *counts, int num_sms, cudaStream_t st)
{
    auto k = fb_profile_mma_kernel<MS, CAP, PHILOX, RESIDENT>;
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    const int64_t ctas = (n * chunks + warps - 1) / warps;
    k<<<fb_grid_for(k, warps * 32, smem, ctas, num_sms), warps * 32, smem, st>>>(net, n, boxes, samples, seed, first_index,
                                                                             nsamp, chunks, chunk_len, counts);
    return cudaGetLastError();
}

// Register tile: networks up to 64 wide use the instantiation that holds 8 tiles per layer, wider ones (up to 160)
// the one that holds 20.  The weights stay in shared memory whenever they fit (every reference model: BM-4 takes 191 KB).
cudaError_t fb_profile_mma(const FbMmaNet &net, int64_t n, const int32_t *boxes, const int32_t *samples, uint64_t seed,
                           int64_t first_index, int nsamp, int chunks, int chunk_len, uint32_t *counts, int num_sms,
                           cudaStream_t st)
{
    if (net.maxnb > 20) return cudaErrorInvalidConfiguration;
    const bool wide = net.maxnb > 8;
    int ms = 1, warps = FB_MMA_WARPS;
    bool resident = fb_mma_smem_bytes(net, ms, warps, true) <= 227 * 1024;
    if (const int v = fb_env_int("FAIRIFY_B200_MMA_MS")) ms = v;
    if (const int v = fb_env_int("FAIRIFY_B200_MMA_WARPS")) warps = v;
    if (fb_env_int("FAIRIFY_B200_MMA_NONRESIDENT")) resident = false;
    if ((ms != 1 && ms != 2) || warps < 1 || warps > FB_MMA_WARPS) return cudaErrorInvalidValue;
    const size_t smem = fb_mma_smem_bytes(net, ms, warps, resident);
    if (smem > 227 * 1024) return cudaErrorInvalidConfiguration;
    const bool philox = samples == nullptr;
#define FB_MMA_GO(M, C, P, R) \
    return fb_mma_launch<M, C, P, R>(net, warps, smem, n, boxes, samples, seed, first_index, nsamp, chunks, chunk_len, counts, num_sms, st)
#define FB_MMA_PR(M, C)                                   \
    if (philox && resident) FB_MMA_GO(M, C, true, true);  \
    if (philox) FB_MMA_GO(M, C, true, false);             \
    if (resident) FB_MMA_GO(M, C, false, true);           \
    FB_MMA_GO(M, C, false, false)
    if (wide) { FB_MMA_PR(1, 20); }          // (two sample blocks of 20 tiles do not fit the register file)
    if (ms == 2) { FB_MMA_PR(2, 8); }
    FB_MMA_PR(1, 8);
}
