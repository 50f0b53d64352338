// fb_common.cuh -- shared device helpers of the fairify_b200 pre-solver kernels (sm_100a).
//
// Network description comes in two flavours:
//   * Spec<D0, D1, ..., 1>: layer widths are template constants.  Weights travel as a __grid_constant__
//     kernel parameter, so every weight reaches its DFMA/DMUL through a uniform register filled from the
//     constant bank (no global or shared load at all) and activations / bounds live in registers.  Used for
//     every architecture in the fb_spec_*.cu lists (all reference models whose parameters fit the 32 KB
//     kernel-parameter space: up to ~3 700 weights).
//   * GenNet / FbTileNet: runtime widths, weights in global memory (staged to shared memory where a kernel
//     reuses them).  Covers everything else (AC-4/5, BM-4/7, unknown nets): the big ones through tiled,
//     matrix-product shaped kernels, the small ones with one sample per lane.
// Both implement exactly the arithmetic of oracle/fairify_oracle.c: term order i ascending, one fused
// multiply-add per term in the forward passes, separately rounded product and sum in the interval
// recurrence, bias added last.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define FB_MAX_LAYERS 16
#define FB_MAXW 160            // widest layer the generic path supports (reference max: 150, BM-4)
#define FB_MAX_NW 32           // mask words per partition (N <= 1024 neurons)
#define FB_WARP 32

// ---------------------------------------------------------------------------------------------------
// Philox4x32-10, counter = (sample, partition_lo, partition_hi, block), key = (seed_lo, seed_hi).
// Bit-identical to philox4x32_10() in the oracle.
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ void fb_philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                 uint32_t k0, uint32_t k1, uint32_t (&out)[4])
{
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        // (written with separate high/low products: measured faster on sm_100a than the 64-bit-product form,
        //  which costs extra register moves in the fully unrolled profile kernel)
#if defined(FB_PHILOX_WIDE) && FB_PHILOX_WIDE
        const uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        const uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0, hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
#else
        const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
#endif
        const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// ---------------------------------------------------------------------------------------------------
// Bounded integers from Philox words.  A feature with `range` values takes the high half of word * range; the
// low half of that product is again a uniform 32-bit word (up to a relative bias of range / 2^32), so it can
// serve the next feature -- the batched form of the multiply-shift method.  How many features share one word
// (G) depends on the partition's widest feature range, so that the product of the ranges drawn from one word
// never exceeds 2^16 and every value's probability stays within 2^-16 (relative) of uniform:
//   widest range <= 16  -> G = 4,   <= 256 -> G = 2,   otherwise G = 1 (one word per feature).
// Feature i uses word i / G of the sample's word stream; word q is output q % 4 of the Philox block with
// counter (sample, partition_lo, partition_hi, q / 4).  Identical in the oracle (fo_sample_row).
// ---------------------------------------------------------------------------------------------------
#define FB_DRAW4_MAX_RANGE 16u
#define FB_DRAW2_MAX_RANGE 256u
__host__ __device__ __forceinline__ int fb_draw_group(uint32_t widest_range)
{
    return widest_range <= FB_DRAW4_MAX_RANGE ? 4 : (widest_range <= FB_DRAW2_MAX_RANGE ? 2 : 1);
}

// one feature: value in [lo, lo + range), and the word handed on to the next feature of the group
__device__ __forceinline__ int32_t fb_draw_next(uint32_t &word, int32_t lo, uint32_t range)
{
    const uint64_t prod = (uint64_t)word * range;
    word = (uint32_t)prod;
    return lo + (int32_t)(uint32_t)(prod >> 32);
}

// Runtime-width form: all D0 features of sample s of partition pidx; box = this partition's (lo, hi) pairs.
// OUT is double or float.
template <typename OUT>
__device__ __forceinline__ void fb_draw_row(const int32_t *__restrict__ box, int D0, int G, uint32_t s, int64_t pidx,
                                            uint64_t seed, OUT *x, int stride = 1)
{
    const int nwords = (D0 + G - 1) / G;
    for (int blk = 0; blk * 4 < nwords; ++blk) {
        uint32_t r[4];
        fb_philox4x32_10(s, (uint32_t)pidx, (uint32_t)((uint64_t)pidx >> 32), (uint32_t)blk, (uint32_t)seed,
                         (uint32_t)(seed >> 32), r);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            uint32_t word = r[k];
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                const int i = (blk * 4 + k) * G + g;
                if (g < G && i < D0) {
                    const int lo = __ldg(box + 2 * i), hi = __ldg(box + 2 * i + 1);
                    x[(int64_t)i * stride] = (OUT)fb_draw_next(word, lo, (uint32_t)(hi - lo + 1));
                }
            }
        }
    }
}

// widest feature range of a box (every lane computes the same value)
__device__ __forceinline__ uint32_t fb_box_widest(const int32_t *__restrict__ box, int D0)
{
    uint32_t widest = 0;
    for (int i = 0; i < D0; ++i) widest = max(widest, (uint32_t)(__ldg(box + 2 * i + 1) - __ldg(box + 2 * i) + 1));
    return widest;
}

// ---------------------------------------------------------------------------------------------------
// Compile-time network shape.
// ---------------------------------------------------------------------------------------------------
template <int... D>
struct Spec {
    static constexpr int NL = (int)sizeof...(D) - 1;     // weight layers
    static constexpr int ND = (int)sizeof...(D);
    __host__ __device__ static constexpr int dim(int l) { constexpr int d[ND] = {D...}; return d[l]; }
    __host__ __device__ static constexpr int woff(int l) { int o = 0; for (int k = 0; k < l; ++k) o += dim(k) * dim(k + 1); return o; }
    __host__ __device__ static constexpr int boff(int l) { int o = 0; for (int k = 0; k < l; ++k) o += dim(k + 1); return o; }
    static constexpr int P = woff(NL);
    static constexpr int N = boff(NL);
    static constexpr int D0 = dim(0);
    static constexpr int NW = (N + 31) / 32;              // mask words per partition
    // Position of weight (input i -> neuron j) of layer l inside SpecParams::w.  Even layers are stored [out][in],
    // odd layers [in][out] (the Keras layout): fb_spec_forward walks an even layer neuron by neuron over ascending
    // inputs and folds each finished neuron into all accumulators of the following odd layer, so in both cases the
    // weights it needs next are neighbours in the constant bank and arrive two per 128-bit uniform load.
    __host__ __device__ static constexpr int widx(int l, int i, int j)
    {
        return (l & 1) ? woff(l) + i * dim(l + 1) + j : woff(l) + j * dim(l) + i;
    }
};

// Kernel parameters are limited to 32 764 bytes.  Where float64 weights + float64 biases would not leave room for the
// other arguments (GC-5: 4004 + 125 values), the biases are kept as the float32 values they are anyway and widened where
// they are added (exact).
#define FB_PARAM_BUDGET (32764 - 224)
template <class S, typename T, bool NARROW_BIAS = (sizeof(T) * (S::P + S::N) > FB_PARAM_BUDGET)>
struct SpecParams {
    T w[S::P];      // layer-major, element order given by S::widx (filled by fb_spec_fill from the Keras layout)
    T b[S::N];
};
template <class S, typename T>
struct SpecParams<S, T, true> {
    T w[S::P];
    float b[S::N];
};

// host: copy a model given layer-major with [in][out] row-major kernels (what Keras get_weights() returns)
template <class S, typename T>
static inline void fb_spec_fill(SpecParams<S, T> &p, const T *W, const T *B)
{
    for (int l = 0; l < S::NL; ++l)
        for (int i = 0; i < S::dim(l); ++i)
            for (int j = 0; j < S::dim(l + 1); ++j) p.w[S::widx(l, i, j)] = W[S::woff(l) + i * S::dim(l + 1) + j];
    for (int k = 0; k < S::N; ++k) p.b[k] = static_cast<decltype(p.b[0] + 0)>(B[k]);   // float32-valued either way: exact
}

// ---------------------------------------------------------------------------------------------------
// Runtime network shape (generic path).  W64/W32 point at device copies with the same layout as above.
// ---------------------------------------------------------------------------------------------------
struct GenNet {
    int L, N, P, maxw;
    int dims[FB_MAX_LAYERS + 1];
    int woff[FB_MAX_LAYERS + 1];
    int boff[FB_MAX_LAYERS + 1];
    const double *W64, *B64;
    const float *W32, *B32;
    const uint32_t *unb;      // [NW] neuron has a non-zero weight on an input its singular-verification query leaves unbounded
};

// ---------------------------------------------------------------------------------------------------
// small arithmetic helpers
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool fb_pos(double v) { return __double_as_longlong(v) > 0; }   // v > 0 (no NaNs here)
// v > 0 from the high word alone: exact unless v is a positive subnormal below 2^-1042.  Only used by the specialised
// profile kernel, which fb_model_create selects only for models that cannot produce a subnormal pre-activation
// (fb_model_subnormal_free in fb_api.cu).
__device__ __forceinline__ bool fb_pos_hi(double v) { return __double2hiint(v) > 0; }
__device__ __forceinline__ bool fb_pos(float v) { return __float_as_int(v) > 0; }
__device__ __forceinline__ bool fb_nonzero(double v) { return (__double_as_longlong(v) << 1) != 0; }
__device__ __forceinline__ bool fb_signbit(double v) { return __double2hiint(v) < 0; }

__device__ __forceinline__ double fb_fma(double a, double b, double c) { return __fma_rn(a, b, c); }
__device__ __forceinline__ float fb_fma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ double fb_add(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ float fb_add(float a, float b) { return __fadd_rn(a, b); }

// K samples carried side by side in one thread (register blocking of the profile kernel: every weight
// fetched from the constant bank feeds K fused multiply-adds)
template <int K>
struct VecD {
    double v[K];
    __device__ __forceinline__ VecD() {}
    __device__ __forceinline__ explicit VecD(double s)
    {
#pragma unroll
        for (int k = 0; k < K; ++k) v[k] = s;
    }
};
template <int K>
__device__ __forceinline__ VecD<K> fb_fma(double w, const VecD<K> &x, const VecD<K> &acc)
{
    VecD<K> r;
#pragma unroll
    for (int k = 0; k < K; ++k) r.v[k] = __fma_rn(w, x.v[k], acc.v[k]);
    return r;
}
template <int K>
__device__ __forceinline__ VecD<K> fb_add(const VecD<K> &a, double b)
{
    VecD<K> r;
#pragma unroll
    for (int k = 0; k < K; ++k) r.v[k] = __dadd_rn(a.v[k], b);
    return r;
}

// Specialised architectures above this many weights: relaxed partition screens run on the tiled runtime-shape kernel and
// the strict ones carry no interval pass (the fully unrolled interval forward takes minutes per kernel to compile)
#ifndef FB_SCREEN_SPEC_RELAXED_MAXP
#define FB_SCREEN_SPEC_RELAXED_MAXP 1500
#endif

// verdict byte
#define FB_VERDICT_CERTIFIED 1u     // output logit provably one-signed over the box (outward-rounded enclosure)
#define FB_VERDICT_CE        2u     // sampled counterexample pair found by the partition screen
#define FB_VERDICT_SLIVER    4u     // some neuron's bound is within rounding error of 0: host Z3 decides s_dead
#define FB_VERDICT_SKIPPED   8u     // the partition screen could not encode this box (protected range too wide)
// partition status codes of the hybrid prediction
#define FB_STATUS_NONE 0
#define FB_STATUS_SAT 1
#define FB_STATUS_UNSAT 2
#define FB_STATUS_UNKNOWN 3
