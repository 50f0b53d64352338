// fb_launch.cuh -- grid sizing shared by every launcher.
#pragma once
#include <algorithm>
#include <cuda_runtime.h>
#include <stdint.h>

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
