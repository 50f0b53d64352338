// fb_spec_ops_fwd.cuh -- host side of one Spec<...>: parameter structs and the launchers of the forward-pass kernels;
// included by the fb_specF_*.cu instantiation units.
#pragma once
#include <cstdlib>
#include <cstring>
#include <new>

#include "fb_launch.cuh"
#include "fb_registry.h"
#include "fb_spec_fwd.cuh"

template <class S>
struct FbSpecFwdOps {
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
                                        int chunk_len, int pa_col, int ra_col, int ra_thr, unsigned long long *ce_key, int num_sms,
                                        cudaStream_t st)
    {
        const P32 &prm = *static_cast<const P32 *>(p32);
        const int64_t ctas = (n * chunks + (FB_PROFILE_THREADS / 32) - 1) / (FB_PROFILE_THREADS / 32);
        auto launch = [&](auto k) {
            k<<<fb_grid_for(k, FB_PROFILE_THREADS, 0, ctas, num_sms), FB_PROFILE_THREADS, 0, st>>>(
                prm, n, boxes, samples, seed, first_index, nsamp, chunks, chunk_len, pa_col, ra_col, ra_thr, ce_key);
        };
        if (ra_col >= 0) {
            // (relaxed queries of the bigger architectures go to the tiled runtime-shape screen, fb_api.cu: their fully
            // unrolled sweep + interval pass would take ten minutes each to compile)
            if constexpr (S::P <= FB_SCREEN_SPEC_RELAXED_MAXP) {
                if (philox) launch(fb_partition_screen_spec_kernel<S, true, true>);
                else launch(fb_partition_screen_spec_kernel<S, false, true>);
            } else {
                return cudaErrorNotSupported;
            }
        } else {
            if (philox) launch(fb_partition_screen_spec_kernel<S, true, false>);
            else launch(fb_partition_screen_spec_kernel<S, false, false>);
        }
        return cudaGetLastError();
    }
};

#define FB_SPEC_SHAPE(...) {Spec<__VA_ARGS__>::ND, {__VA_ARGS__}}
#define FB_SPEC_FWD_ENTRY(...)                                                                                        \
    {FB_SPEC_SHAPE(__VA_ARGS__), &FbSpecFwdOps<Spec<__VA_ARGS__>>::make_params64,                                     \
     &FbSpecFwdOps<Spec<__VA_ARGS__>>::make_params32, &FbSpecFwdOps<Spec<__VA_ARGS__>>::free_params,                  \
     &FbSpecFwdOps<Spec<__VA_ARGS__>>::forward64, &FbSpecFwdOps<Spec<__VA_ARGS__>>::forward32,                        \
     &FbSpecFwdOps<Spec<__VA_ARGS__>>::pair_screen, &FbSpecFwdOps<Spec<__VA_ARGS__>>::partition_screen},
