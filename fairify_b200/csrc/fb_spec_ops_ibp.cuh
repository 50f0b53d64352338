// fb_spec_ops_ibp.cuh -- launcher of the specialised bounds kernel for one Spec<...>; included by the fb_specI_*.cu
// instantiation units.
#pragma once
#include <cstdlib>

#include "fb_launch.cuh"
#include "fb_registry.h"
#include "fb_spec_ibp.cuh"

// The unrolled one-thread-per-partition kernel holds four bound arrays per layer in registers and spills beyond
// ~1200 weights; from there the warp-per-partition runtime-shape kernel is faster (measured on B200, 10^6 boxes:
// BM-2 1040 weights 2.2 vs 3.6 ms, AC-2 1400: 8.2 vs 7.4, BM-1 2064: 11.8 vs 5.9, AC-7 3556: 19.7 vs 8.7), so
// larger shapes get no specialised bounds kernel at all (their table row holds a null launcher).
#ifndef FB_IBP_SPEC_MAX_WEIGHTS
#define FB_IBP_SPEC_MAX_WEIGHTS 1200
#endif

typedef cudaError_t (*FbIbpLaunch)(const void *p64, int64_t n, const int32_t *boxes, const uint32_t *counts,
                                   const double *tol, const uint32_t *unb, const FbPruneOut *out, int num_sms,
                                   cudaStream_t st);

template <class S, bool BUILD = (S::P <= FB_IBP_SPEC_MAX_WEIGHTS || fb_ibp_fused2_ok<S>())>
struct FbSpecIbpOps {
    using P64 = SpecParams<S, double>;
    static cudaError_t ibp(const void *p64, int64_t n, const int32_t *boxes, const uint32_t *counts, const double *tol,
                           const uint32_t *unb, const FbPruneOut *out, int num_sms, cudaStream_t st)
    {
        const P64 &prm = *static_cast<const P64 *>(p64);
        // FAIRIFY_B200_IBP_UNROLLED=1 keeps the fully unrolled kernel for the shapes the looped one covers (comparisons)
        static const bool unrolled = [] { const char *v = getenv("FAIRIFY_B200_IBP_UNROLLED"); return v && v[0] == '1'; }();
        if constexpr (fb_ibp_loop_ok<S>()) {
            if (!unrolled || S::P > FB_IBP_SPEC_MAX_WEIGHTS) {
                auto k = fb_ibp_loop_kernel<S>;
                const size_t smem = fb_ibp_fused2_ok<S>() ? 0 : sizeof(double) * 4 * FB_IBP_LOOP_THREADS * fb_ibp_max_hidden<S>();
                cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
                if (e != cudaSuccess) return e;
                const int64_t ctas = (n + FB_IBP_LOOP_THREADS - 1) / FB_IBP_LOOP_THREADS;
                k<<<fb_grid_for(k, FB_IBP_LOOP_THREADS, smem, ctas, num_sms), FB_IBP_LOOP_THREADS, smem, st>>>(prm, n, boxes, counts,
                                                                                                          tol, unb, *out);
                return cudaGetLastError();
            }
        }
        if constexpr (S::P <= FB_IBP_SPEC_MAX_WEIGHTS) {
            auto k = fb_ibp_spec_kernel<S>;
            const int64_t ctas = (n + FB_IBP_THREADS - 1) / FB_IBP_THREADS;
            k<<<fb_grid_for(k, FB_IBP_THREADS, 0, ctas, num_sms), FB_IBP_THREADS, 0, st>>>(prm, n, boxes, counts, tol, unb, *out);
            return cudaGetLastError();
        } else {
            return cudaErrorNotSupported;          // (no unrolled kernel is built for shapes this large)
        }
    }

    static constexpr FbIbpLaunch launcher = &FbSpecIbpOps::ibp;
};
template <class S>
struct FbSpecIbpOps<S, false> {
    static constexpr FbIbpLaunch launcher = nullptr;
};

#define FB_SPEC_SHAPE(...) {Spec<__VA_ARGS__>::ND, {__VA_ARGS__}}
#define FB_SPEC_IBP_ENTRY(...) {FB_SPEC_SHAPE(__VA_ARGS__), FbSpecIbpOps<Spec<__VA_ARGS__>>::launcher},
