// fb_spec_ops_profile.cuh -- launcher of the specialised activation-profile kernel for one Spec<...>; included by the
// fb_specP_*.cu instantiation units.
#pragma once
#include <cstdlib>

#include "fb_launch.cuh"
#include "fb_registry.h"
#include "fb_spec_profile.cuh"

template <class S>
struct FbSpecProfileOps {
    using P64 = SpecParams<S, double>;

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

};

#define FB_SPEC_SHAPE(...) {Spec<__VA_ARGS__>::ND, {__VA_ARGS__}}
#define FB_SPEC_PROFILE_ENTRY(...) {FB_SPEC_SHAPE(__VA_ARGS__), &FbSpecProfileOps<Spec<__VA_ARGS__>>::profile},
