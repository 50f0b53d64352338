// fb_registry.h -- host-side table of specialised kernel launchers (one row per architecture in
// FB_SPEC_LIST).  The C-ABI layer (fb_api.cu) looks a model's layer widths up here at fb_model_create
// time and falls back to the runtime-shape kernels when there is no row.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "fb_common.cuh"
#include "fb_spec.cuh"

struct FbLaunchGeom {
    int num_sms;
};

struct FbSpecEntry {
    int nd;                                   // number of widths (layers + 1)
    int dims[FB_MAX_LAYERS + 1];
    void *(*make_params64)(const double *W, const double *B);
    void *(*make_params32)(const float *W, const float *B);
    void (*free_params)(void *p64, void *p32);
    cudaError_t (*profile)(const void *p64, bool philox, int64_t n, const int32_t *boxes, const int32_t *samples,
                           uint64_t seed, int64_t first_index, int nsamp, int chunks, int chunk_len, uint32_t *counts,
                           int num_sms, cudaStream_t st);
    cudaError_t (*ibp)(const void *p64, int64_t n, const int32_t *boxes, const uint32_t *counts, const double *tol,
                       const FbPruneOut *out, int num_sms, cudaStream_t st);
    cudaError_t (*forward64)(const void *p64, int64_t m, const double *X, const uint32_t *dead_bits, double *logits,
                             int num_sms, cudaStream_t st);
    cudaError_t (*forward32)(const void *p32, int64_t m, const float *X, const uint32_t *dead_bits, float *logits,
                             int num_sms, cudaStream_t st);
    cudaError_t (*pair_screen)(const void *p32, int64_t m, const float *X, const FbPaCols *pa, int k_alts,
                               const float *alts, uint8_t *orig_pred, int32_t *first_flip, int num_sms, cudaStream_t st);
    cudaError_t (*partition_screen)(const void *p32, bool philox, int64_t n, const int32_t *boxes,
                                    const int32_t *samples, uint64_t seed, int64_t first_index, int nsamp, int chunks,
                                    int chunk_len, int pa_col, unsigned long long *ce_key, int num_sms, cudaStream_t st);
};

struct FbSpecTable {
    const FbSpecEntry *entries;
    int count;
};

// defined in fb_spec_*.cu
extern const FbSpecTable fb_spec_table_a, fb_spec_table_b, fb_spec_table_c, fb_spec_table_d, fb_spec_table_e, fb_spec_table_f,
    fb_spec_table_g;
