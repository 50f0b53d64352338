// fb_spec_ops_mma.cuh -- launcher of the shape-specialised DMMA profile kernel for one Spec<...>; included by the
// fb_specM_*.cu instantiation units.
#pragma once
#include <cstdlib>

#include "fb_launch.cuh"
#include "fb_mma_spec.cuh"
#include "fb_registry.h"

template <class S>
struct FbSpecMmaOps {
    template <int MS, bool PHILOX>
    static cudaError_t go(const FbMmaNet &net, int warps, int64_t n, const int32_t *boxes, const int32_t *samples,
                          uint64_t seed, int64_t first_index, int nsamp, int chunks, int chunk_len, uint32_t *counts,
                          int num_sms, cudaStream_t st)
    {
        auto k = fb_profile_mma_spec_kernel<S, MS, PHILOX>;
        const size_t smem = fb_mma_smem_bytes(net, MS, warps, true);
        if (smem > 227 * 1024) return cudaErrorInvalidConfiguration;
        cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        const int64_t ctas = (n * chunks + warps - 1) / warps;
        k<<<fb_grid_for(k, warps * 32, smem, ctas, num_sms), warps * 32, smem, st>>>(net, n, boxes, samples, seed, first_index,
                                                                                 nsamp, chunks, chunk_len, counts);
        return cudaGetLastError();
    }

    // MS = sample blocks (of 8) per warp step.  The instantiation unit says which values it builds
    // (FB_MMA_SPEC_MS_MASK: bit 0 = 1, bit 1 = 2, bit 2 = 4 blocks) and which one is the default;
    // FAIRIFY_B200_MMA_MS / FAIRIFY_B200_MMA_WARPS override (tuning aids)
    template <int MS>
    static cudaError_t go_ms(bool philox, const FbMmaNet &net, int warps, int64_t n, const int32_t *boxes, const int32_t *samples,
                             uint64_t seed, int64_t first_index, int nsamp, int chunks, int chunk_len, uint32_t *counts,
                             int num_sms, cudaStream_t st)
    {
        if (philox) return go<MS, true>(net, warps, n, boxes, samples, seed, first_index, nsamp, chunks, chunk_len, counts, num_sms, st);
        return go<MS, false>(net, warps, n, boxes, samples, seed, first_index, nsamp, chunks, chunk_len, counts, num_sms, st);
    }

    static cudaError_t profile(const FbMmaNet &net, int64_t n, const int32_t *boxes, const int32_t *samples, uint64_t seed,
                               int64_t first_index, int nsamp, int chunks, int chunk_len, uint32_t *counts, int num_sms,
                               cudaStream_t st)
    {
        int ms = FB_MMA_SPEC_DEFAULT_MS, warps = FB_MMA_WARPS;
        if (const char *v = getenv("FAIRIFY_B200_MMA_MS")) ms = atoi(v);
        if (const char *v = getenv("FAIRIFY_B200_MMA_WARPS")) warps = atoi(v);
        if (warps < 1 || warps > FB_MMA_WARPS) return cudaErrorInvalidValue;
        const bool philox = samples == nullptr;
#if FB_MMA_SPEC_MS_MASK & 4
        if (ms == 4) return go_ms<4>(philox, net, warps, n, boxes, samples, seed, first_index, nsamp, chunks, chunk_len, counts, num_sms, st);
#endif
#if FB_MMA_SPEC_MS_MASK & 2
        if (ms == 2) return go_ms<2>(philox, net, warps, n, boxes, samples, seed, first_index, nsamp, chunks, chunk_len, counts, num_sms, st);
#endif
        return go_ms<(FB_MMA_SPEC_MS_MASK & 1) ? 1 : FB_MMA_SPEC_DEFAULT_MS>(philox, net, warps, n, boxes, samples, seed, first_index, nsamp, chunks, chunk_len, counts, num_sms, st);
    }
};

#define FB_SPEC_SHAPE(...) {Spec<__VA_ARGS__>::ND, {__VA_ARGS__}}
#define FB_SPEC_MMA_ENTRY(...) {FB_SPEC_SHAPE(__VA_ARGS__), &FbSpecMmaOps<Spec<__VA_ARGS__>>::profile},
