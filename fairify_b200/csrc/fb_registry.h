// fb_registry.h -- host-side tables of specialised kernel launchers, one row per architecture in fb_spec_lists.h and
// one table per kernel family, so that each family is its own set of compile units (the fully unrolled kernels are
// slow to build).  The C-ABI layer (fb_api.cu) looks a model's layer widths up here at fb_model_create time and falls
// back to the runtime-shape kernels when there is no row.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "fb_common.cuh"

struct FbPruneOut;
struct FbPaCols;

struct FbSpecShape {
    int nd;                                   // number of widths (layers + 1)
    int dims[FB_MAX_LAYERS + 1];
};

// parameter structs + the float/double forward kernels (logits, pair screen, partition screen)
struct FbSpecFwdEntry {
    FbSpecShape shape;
    void *(*make_params64)(const double *W, const double *B);
    void *(*make_params32)(const float *W, const float *B);
    void (*free_params)(void *p64, void *p32);
    cudaError_t (*forward64)(const void *p64, int64_t m, const double *X, const uint32_t *dead_bits, double *logits,
                             int num_sms, cudaStream_t st);
    cudaError_t (*forward32)(const void *p32, int64_t m, const float *X, const uint32_t *dead_bits, float *logits,
                             int num_sms, cudaStream_t st);
    cudaError_t (*pair_screen)(const void *p32, int64_t m, const float *X, const FbPaCols *pa, int k_alts,
                               const float *alts, uint8_t *orig_pred, int32_t *first_flip, int num_sms, cudaStream_t st);
    cudaError_t (*partition_screen)(const void *p32, bool philox, int64_t n, const int32_t *boxes,
                                    const int32_t *samples, uint64_t seed, int64_t first_index, int nsamp, int chunks,
                                    int chunk_len, int pa_col, int ra_col, int ra_thr, unsigned long long *ce_key, int num_sms,
                                    cudaStream_t st);
};

struct FbSpecProfileEntry {
    FbSpecShape shape;
    cudaError_t (*profile)(const void *p64, bool philox, int64_t n, const int32_t *boxes, const int32_t *samples,
                           uint64_t seed, int64_t first_index, int nsamp, int chunks, int chunk_len, uint32_t *counts,
                           int num_sms, cudaStream_t st);
};

// ibp == nullptr: the shape is left to the runtime-shape bounds kernel (see fb_spec_ops_ibp.cuh)
struct FbSpecIbpEntry {
    FbSpecShape shape;
    cudaError_t (*ibp)(const void *p64, int64_t n, const int32_t *boxes, const uint32_t *counts, const double *tol,
                       const uint32_t *unb, const FbPruneOut *out, int num_sms, cudaStream_t st);
};

// shape-specialised DMMA profile kernel (fb_mma_spec.cuh): the wide networks
struct FbMmaNet;
struct FbSpecMmaEntry {
    FbSpecShape shape;
    cudaError_t (*profile)(const FbMmaNet &net, int64_t n, const int32_t *boxes, const int32_t *samples, uint64_t seed,
                           int64_t first_index, int nsamp, int chunks, int chunk_len, uint32_t *counts, int num_sms,
                           cudaStream_t st);
};

template <class E>
struct FbSpecTableT {
    const E *entries;
    int count;
};
typedef FbSpecTableT<FbSpecFwdEntry> FbSpecFwdTable;
typedef FbSpecTableT<FbSpecProfileEntry> FbSpecProfileTable;
typedef FbSpecTableT<FbSpecIbpEntry> FbSpecIbpTable;
typedef FbSpecTableT<FbSpecMmaEntry> FbSpecMmaTable;
extern const FbSpecMmaTable fb_spec_mma_table_a, fb_spec_mma_table_b, fb_spec_mma_table_c, fb_spec_mma_table_d;

// defined in fb_specF_*.cu / fb_specP_*.cu / fb_specI_*.cu
#define FB_SPEC_DECLARE_TABLES(U)                              \
    extern const FbSpecFwdTable fb_spec_fwd_table_##U;         \
    extern const FbSpecProfileTable fb_spec_profile_table_##U; \
    extern const FbSpecIbpTable fb_spec_ibp_table_##U;
FB_SPEC_DECLARE_TABLES(a) FB_SPEC_DECLARE_TABLES(b) FB_SPEC_DECLARE_TABLES(c) FB_SPEC_DECLARE_TABLES(d)
FB_SPEC_DECLARE_TABLES(e) FB_SPEC_DECLARE_TABLES(f) FB_SPEC_DECLARE_TABLES(g)
#define FB_SPEC_FOR_EACH_UNIT(X) X(a) X(b) X(c) X(d) X(e) X(f) X(g)
