// fb_spec_ops.cuh -- host launchers for one Spec<...>; included by the fb_spec_*.cu instantiation units.
#pragma once
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <new>

#include "fb_registry.h"

// grid = enough CTAs to cover the work, capped at what is co-resident (occupancy x SM count) so the
// grid-stride loops run as one balanced wave.
template <class K>
static inline int fb_grid_for(K kernel, int threads, size_t smem, int64_t ctas_needed, int num_sms)
{
    int per_sm = 1;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem) != cudaSuccess || per_sm < 1) per_sm = 1;
    const int64_t cap = (int64_t)per_sm * num_sms;
    return (int)std::max<int64_t>(1, std::min<int64_t>(ctas_needed, cap));
}

template <class S>
struct FbSpecOps {
    using P64 = SpecParams<S, double>;
    using P32 = SpecParams<S, float>;
    static_assert(sizeof(P64) <= FB_PARAM_BUDGET, "parameter struct exceeds the kernel-parameter limit");

    static void *make_params64(const double *W, const double *B)
    {
        P64 *p = new (std::nothrow) P64;
        if (!p) return nullptr;
        fb_spec_fill(*p, W, B);
        return p;
    }
    static void *make_params32(const float *W, const float *B)
    {
        P32 *p = new (std::nothrow) P32;
        if (!p) return nullptr;
        fb_spec_fill(*p, W, B);
        return p;
    }
    static void free_params(void *a, void *b) { delete static_cast<P64 *>(a); delete static_cast<P32 *>(b); }

    static cudaError_t profile(const void *p64, bool philox, int64_t n, const int32_t *boxes, const int32_t *samples,
                               uint64_t seed, int64_t first_index, int nsamp, int chunks, int chunk_len, uint32_t *counts,
                               int num_sms, cudaStream_t st)
    {
        const P64 &prm = *static_cast<const P64 *>(p64);
        constexpr int WPC = FB_PROFILE_THREADS / 32;
        const int64_t units = n * chunks;
        // Units are handed to warps in blocks of `ublock` (a launch's last round leaves some warps one block short:
        // keep that under ~1/64 of a warp's work), 32 at most -- the lanes of a warp classify one unit each
        int ublock = 1;
        while (ublock < 32 && units >= (int64_t)64 * (2 * ublock) * WPC * num_sms) ublock *= 2;
        if (const char *v = getenv("FAIRIFY_B200_UBLOCK")) {          // test hook: force the block size (1 ... 32)
            const int x = atoi(v);
            if (x >= 1 && x <= 32) ublock = x;
        }
        const int64_t ctas = ((units + ublock - 1) / ublock + WPC - 1) / WPC;
        if (philox) {
            // one launch per draw class; each takes the partitions of its class (see the kernel)
            auto k4 = fb_profile_spec_kernel<S, true, 4>;
            auto k2 = fb_profile_spec_kernel<S, true, 2>;
            auto k1 = fb_profile_spec_kernel<S, true, 1>;
            const int grid = fb_grid_for(k4, FB_PROFILE_THREADS, 0, ctas, num_sms);
            k4<<<grid, FB_PROFILE_THREADS, 0, st>>>(prm, n, boxes, samples, seed, first_index, nsamp, chunks, chunk_len, ublock, counts);
            k2<<<grid, FB_PROFILE_THREADS, 0, st>>>(prm, n, boxes, samples, seed, first_index, nsamp, chunks, chunk_len, ublock, counts);
            k1<<<grid, FB_PROFILE_THREADS, 0, st>>>(prm, n, boxes, samples, seed, first_index, nsamp, chunks, chunk_len, ublock, counts);
        } else {
            auto k = fb_profile_spec_kernel<S, false, 1>;
            k<<<fb_grid_for(k, FB_PROFILE_THREADS, 0, ctas, num_sms), FB_PROFILE_THREADS, 0, st>>>(
                prm, n, boxes, samples, seed, first_index, nsamp, chunks, chunk_len, ublock, counts);
        }
        return cudaGetLastError();
    }

    static cudaError_t ibp(const void *p64, int64_t n, const int32_t *boxes, const uint32_t *counts, const double *tol,
                           const FbPruneOut *out, int num_sms, cudaStream_t st)
    {
        const P64 &prm = *static_cast<const P64 *>(p64);
        auto k = fb_ibp_spec_kernel<S>;
        const int64_t ctas = (n + FB_IBP_THREADS - 1) / FB_IBP_THREADS;
        k<<<fb_grid_for(k, FB_IBP_THREADS, 0, ctas, num_sms), FB_IBP_THREADS, 0, st>>>(prm, n, boxes, counts, tol, *out);
        return cudaGetLastError();
    }

    static cudaError_t forward64(const void *p64, int64_t m, const double *X, const uint32_t *dead_bits, double *logits,
                                 int num_sms, cudaStream_t st)
    {
        const P64 &prm = *static_cast<const P64 *>(p64);
        auto k = fb_forward_spec_kernel<S, double>;
        const int64_t ctas = (m + FB_FWD_THREADS - 1) / FB_FWD_THREADS;
        k<<<fb_grid_for(k, FB_FWD_THREADS, 0, ctas, num_sms), FB_FWD_THREADS, 0, st>>>(prm, m, X, dead_bits, logits);
        return cudaGetLastError();
    }

    static cudaError_t forward32(const void *p32, int64_t m, const float *X, const uint32_t *dead_bits, float *logits,
                                 int num_sms, cudaStream_t st)
    {
        const P32 &prm = *static_cast<const P32 *>(p32);
        auto k = fb_forward_spec_kernel<S, float>;
        const int64_t ctas = (m + FB_FWD_THREADS - 1) / FB_FWD_THREADS;
        k<<<fb_grid_for(k, FB_FWD_THREADS, 0, ctas, num_sms), FB_FWD_THREADS, 0, st>>>(prm, m, X, dead_bits, logits);
        return cudaGetLastError();
    }

    static cudaError_t pair_screen(const void *p32, int64_t m, const float *X, const FbPaCols *pa, int k_alts,
                                   const float *alts, uint8_t *orig_pred, int32_t *first_flip, int num_sms, cudaStream_t st)
    {
        const P32 &prm = *static_cast<const P32 *>(p32);
        auto k = fb_pair_screen_spec_kernel<S>;
        const int64_t ctas = (m + FB_FWD_THREADS - 1) / FB_FWD_THREADS;
        k<<<fb_grid_for(k, FB_FWD_THREADS, 0, ctas, num_sms), FB_FWD_THREADS, 0, st>>>(prm, m, X, *pa, k_alts, alts,
                                                                                     orig_pred, first_flip);
        return cudaGetLastError();
    }

    static cudaError_t partition_screen(const void *p32, bool philox, int64_t n, const int32_t *boxes,
                                        const int32_t *samples, uint64_t seed, int64_t first_index, int nsamp, int chunks,
                                        int chunk_len, int pa_col, unsigned long long *ce_key, int num_sms, cudaStream_t st)
    {
        const P32 &prm = *static_cast<const P32 *>(p32);
        const int64_t ctas = (n * chunks + (FB_PROFILE_THREADS / 32) - 1) / (FB_PROFILE_THREADS / 32);
        if (philox) {
            auto k = fb_partition_screen_spec_kernel<S, true>;
            k<<<fb_grid_for(k, FB_PROFILE_THREADS, 0, ctas, num_sms), FB_PROFILE_THREADS, 0, st>>>(
                prm, n, boxes, samples, seed, first_index, nsamp, chunks, chunk_len, pa_col, ce_key);
        } else {
            auto k = fb_partition_screen_spec_kernel<S, false>;
            k<<<fb_grid_for(k, FB_PROFILE_THREADS, 0, ctas, num_sms), FB_PROFILE_THREADS, 0, st>>>(
                prm, n, boxes, samples, seed, first_index, nsamp, chunks, chunk_len, pa_col, ce_key);
        }
        return cudaGetLastError();
    }
};

#define FB_SPEC_ENTRY(...)                                                                                           \
    {Spec<__VA_ARGS__>::ND, {__VA_ARGS__}, &FbSpecOps<Spec<__VA_ARGS__>>::make_params64,                             \
     &FbSpecOps<Spec<__VA_ARGS__>>::make_params32, &FbSpecOps<Spec<__VA_ARGS__>>::free_params,                       \
     &FbSpecOps<Spec<__VA_ARGS__>>::profile, &FbSpecOps<Spec<__VA_ARGS__>>::ibp,                                     \
     &FbSpecOps<Spec<__VA_ARGS__>>::forward64, &FbSpecOps<Spec<__VA_ARGS__>>::forward32,                             \
     &FbSpecOps<Spec<__VA_ARGS__>>::pair_screen, &FbSpecOps<Spec<__VA_ARGS__>>::partition_screen}
