// fb_spec_fwd.cuh -- compile-time-specialised forward pass and the kernels built on it alone: logits, pair screen,
// partition screen (see fb_common.cuh).
//
// Every loop over neurons / inputs is fully unrolled; weights are immediate constant-bank operands
// (the parameter struct is __grid_constant__), activations are registers.
#pragma once
#include "fb_common.cuh"

// ---------------------------------------------------------------------------------------------------
// forward pass, one sample per thread.  reference: layer_net / net, utils/AC-1-Model-Functions.py:16-32
// Visit::neuron(idx, last, pre) returns the activation fed to the next layer; idx and last fold to
// constants because every call site is inside fully unrolled loops.
// ---------------------------------------------------------------------------------------------------
//
// Layers are evaluated in pairs so that a wide hidden layer never sits in registers: layer l goes neuron by
// neuron (terms i ascending), and each finished activation h_j is folded at once into the accumulators of
// layer l+1 (acc2[k] += w2[j][k] * h_j), whose own terms thereby also arrive for j ascending.  Every neuron
// of either layer still sees one fused multiply-add per term in ascending input order and its bias last, so
// the results are bit-identical to the plain layer-by-layer evaluation (and to the oracle).  Live state is the
// inputs of layer l plus the accumulators of layer l+1 -- e.g. 13 + 1 values for a 13-100-1 network instead
// of 13 + 100.
#ifndef FB_SPEC_JB
#define FB_SPEC_JB 1            // neurons evaluated side by side when a thread carries two samples
#endif
// ... and when it carries one (then a neuron's dot product is the only chain in flight): 8 for layers of 64 neurons or
// more, else 4.  Measured on B200, % of FP64 peak for blocks of 1 / 2 / 4 / 8: BM-1 (16-64-16-1) 38 / 39 / 42 / 51,
// BM-8 42 / 41 / 48 / 48, AC-7 49 / 49 / 51 / 51, BM-2 (16-32-16-1) 58 / 62 / 61 / 55.
// Per-shape choices that only change ptxas' schedule, taken from measurements on B200 (tools/time_profile.py, % of
// FP64 peak): neuron block for two-sample shapes, and whether the box waits in shared memory for shapes above 64 neurons.
//   13-100-1 (AC-2): block 1 / 2 / 4 = 63 / 64 / 71, shared-memory box 63     20-100-1 (GC-2): 63 / 62 / 53, box 72
//   16-100-1 (BM-3): 70 / 69 / 70, box 63
template <class S> struct FbTune { static constexpr int jb2 = FB_SPEC_JB; static constexpr bool smem_box_wide = false; };
template <> struct FbTune<Spec<13, 100, 1>> { static constexpr int jb2 = 4; static constexpr bool smem_box_wide = false; };
template <> struct FbTune<Spec<20, 100, 1>> { static constexpr int jb2 = FB_SPEC_JB; static constexpr bool smem_box_wide = true; };
template <typename T> struct FbOneSample { static constexpr bool value = false; };
template <> struct FbOneSample<VecD<1>> { static constexpr bool value = true; };
// (the row-per-thread float kernels -- screens, logits -- measured the same with and without blocks)
template <class S, int l, typename T, class PRM, class Visit>
__device__ __forceinline__ T fb_spec_forward(const PRM &prm, const T (&x)[S::dim(l)], Visit &v)
{
    constexpr int IN = S::dim(l), M = S::dim(l + 1), BO = S::boff(l);
    constexpr bool has_next = (l + 1 < S::NL);
    constexpr int L2 = has_next ? l + 1 : l;                      // (kept in range when there is no next layer)
    constexpr int M2 = has_next ? S::dim(L2 + 1) : 1, BO2 = S::boff(L2);
    T acc2[M2];
#pragma unroll
    for (int k = 0; k < M2; ++k) acc2[k] = T(0.0);
    T h = T(0.0);
    // neurons of layer l in blocks of JB (independent dot products side by side).  With two samples per thread ptxas
    // interleaves neighbouring neurons' chains by itself (1, 2, 4, 8 measured the same on B200); with one sample per
    // thread it leaves each dot product a serial chain of ~20-cycle multiply-adds unless the source spells the block out.
    constexpr int JB = FbOneSample<T>::value ? (M >= 64 ? 8 : 4) : FbTune<S>::jb2;
#pragma unroll
    for (int j0 = 0; j0 < M; j0 += JB) {
        T acc[JB];
#pragma unroll
        for (int jj = 0; jj < JB; ++jj) acc[jj] = T(0.0);
#pragma unroll
        for (int i = 0; i < IN; ++i) {
#pragma unroll
            for (int jj = 0; jj < JB; ++jj)
                if (j0 + jj < M) acc[jj] = fb_fma(prm.w[S::widx(l, i, j0 + jj)], x[i], acc[jj]);
        }
#pragma unroll
        for (int jj = 0; jj < JB; ++jj) {
            if (j0 + jj < M) {
                h = v.neuron(BO + j0 + jj, !has_next, fb_add(acc[jj], prm.b[BO + j0 + jj]));
                if constexpr (has_next) {
#pragma unroll
                    for (int k = 0; k < M2; ++k) acc2[k] = fb_fma(prm.w[S::widx(L2, j0 + jj, k)], h, acc2[k]);
                }
            }
        }
    }
    if constexpr (!has_next) return h;                            // the network's single logit (M == 1)
    else {
        constexpr bool last2 = (l + 2 == S::NL);
        T y2[M2];
#pragma unroll
        for (int k = 0; k < M2; ++k) y2[k] = v.neuron(BO2 + k, last2, fb_add(acc2[k], prm.b[BO2 + k]));
        if constexpr (last2) return y2[0];
        else return fb_spec_forward<S, l + 2, T, PRM, Visit>(prm, y2, v);
    }
}

// Interval of float32 values for the screens' one-pass enclosure.  fb_fma / fb_add apply the forward pass's own
// round-to-nearest operation to both ends (rounding is monotone: for x in [x.lo, x.hi] and acc in [acc.lo, acc.hi] the
// computed fma(w, x, acc) lies between the two results), so fb_spec_forward on FbIv encloses every value the float
// forward pass computes for inputs inside the input intervals -- computed values, not real-arithmetic ones.
struct FbIv {
    float lo, hi;
    __device__ __forceinline__ FbIv() {}
    __device__ __forceinline__ explicit FbIv(double v) : lo((float)v), hi((float)v) {}
};
__device__ __forceinline__ FbIv fb_fma(float w, const FbIv &x, const FbIv &acc)
{
    FbIv r;
    const bool up = w >= 0.0f;
    r.lo = __fmaf_rn(w, up ? x.lo : x.hi, acc.lo);
    r.hi = __fmaf_rn(w, up ? x.hi : x.lo, acc.hi);
    return r;
}
__device__ __forceinline__ FbIv fb_add(const FbIv &a, float b)
{
    FbIv r;
    r.lo = __fadd_rn(a.lo, b); r.hi = __fadd_rn(a.hi, b);
    return r;
}
struct FbIvReluVisit {
    __device__ __forceinline__ FbIv neuron(int, bool last, const FbIv &pre)
    {
        if (last) return pre;
        FbIv r;
        r.lo = fb_pos(pre.lo) ? pre.lo : 0.0f; r.hi = fb_pos(pre.hi) ? pre.hi : 0.0f;
        return r;
    }
};

// plain ReLU visitor (logit only)
template <typename T>
struct FbReluVisit {
    __device__ __forceinline__ T neuron(int, bool last, T pre) { return last ? pre : (fb_pos(pre) ? pre : (T)0); }
};

// ReLU + "these neurons are pruned" visitor (dead activations forced to 0 == reference prune_neurons,
// utils/prune.py:950-977, which deletes the column and the next layer's row)
template <typename T, int NW>
struct FbMaskedVisit {
    uint32_t dead[NW];
    __device__ __forceinline__ T neuron(int idx, bool last, T pre)
    {
        if (last) return pre;
        const bool keep = fb_pos(pre) && !((dead[idx >> 5] >> (idx & 31)) & 1u);
        return keep ? pre : (T)0;
    }
};


#ifndef FB_PROFILE_THREADS
#define FB_PROFILE_THREADS 512
#endif

// Compile-time-width form of fb_draw_row: features of the K samples s, s+32, ... of partition pidx, G features per
// Philox word (G = fb_draw_group of the box, warp-uniform).
template <int K>
__device__ __forceinline__ void fb_draw_put(VecD<K> &x, int k, int32_t val) { x.v[k] = (double)val; }
__device__ __forceinline__ void fb_draw_put(float &x, int, int32_t val) { x = (float)val; }

template <int D0, int K, int G, class X>
__device__ __forceinline__ void fb_draw_spec(const int32_t (&lo)[D0], const uint32_t (&range)[D0], int s, int64_t pidx,
                                             uint64_t seed, X (&x)[D0])
{
    constexpr int NWORDS = (D0 + G - 1) / G;
    uint32_t r[K][4];
#pragma unroll
    for (int q = 0; q < NWORDS; ++q) {
        if ((q & 3) == 0) {
#pragma unroll
            for (int k = 0; k < K; ++k)
                fb_philox4x32_10((uint32_t)(s + 32 * k), (uint32_t)pidx, (uint32_t)((uint64_t)pidx >> 32),
                                 (uint32_t)(q >> 2), (uint32_t)seed, (uint32_t)(seed >> 32), r[k]);
        }
#pragma unroll
        for (int k = 0; k < K; ++k) {
            uint32_t word = r[k][q & 3];
#pragma unroll
            for (int g = 0; g < G; ++g)
                if (q * G + g < D0) fb_draw_put(x[q * G + g], k, fb_draw_next(word, lo[q * G + g], range[q * G + g]));
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// logits for a batch of rows, one row per thread (get_y_pred / C-check, utils/verif_utils.py:1040-1047,
// src/AC/Verify-AC.py:229-249).  T = double for integer / float64 rows, float for float32 rows.
// ---------------------------------------------------------------------------------------------------
#define FB_FWD_THREADS 128

template <class S, typename T>
__global__ void __launch_bounds__(FB_FWD_THREADS)
fb_forward_spec_kernel(const __grid_constant__ SpecParams<S, T> prm, int64_t m, const T *__restrict__ X,
                       const uint32_t *__restrict__ dead_bits, T *__restrict__ logits)
{
    constexpr int D0 = S::D0;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < m; r += stride) {
        T x[D0];
#pragma unroll
        for (int i = 0; i < D0; ++i) x[i] = X[r * D0 + i];
        FbMaskedVisit<T, S::NW> vis;
#pragma unroll
        for (int k = 0; k < S::NW; ++k) vis.dead[k] = dead_bits ? dead_bits[k] : 0u;
        logits[r] = fb_spec_forward<S, 0, T>(prm, x, vis);
    }
}

// ---------------------------------------------------------------------------------------------------
// prediction side of CausalDiscriminationDetector.causal_discrimination, src/AC/metrics.py:130-159:
// per base row the original class and the first alternative protected assignment whose class differs.
// ---------------------------------------------------------------------------------------------------
#define FB_MAX_PA 4

struct FbPaCols { int n; int col[FB_MAX_PA]; };

template <class S>
__global__ void __launch_bounds__(FB_FWD_THREADS)
fb_pair_screen_spec_kernel(const __grid_constant__ SpecParams<S, float> prm, int64_t m, const float *__restrict__ X,
                           const FbPaCols pa, int k_alts, const float *__restrict__ alts,
                           uint8_t *__restrict__ orig_pred, int32_t *__restrict__ first_flip)
{
    constexpr int D0 = S::D0;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < m; r += stride) {
        float x0[D0];
#pragma unroll
        for (int i = 0; i < D0; ++i) x0[i] = X[r * D0 + i];
        FbReluVisit<float> vis;
        const bool p0 = fb_pos(fb_spec_forward<S, 0, float>(prm, x0, vis));
        int flip = -1;
        for (int a = 0; a < k_alts && flip < 0; ++a) {
            float x[D0];
#pragma unroll
            for (int i = 0; i < D0; ++i) x[i] = x0[i];
            bool same = true;
            for (int q = 0; q < pa.n; ++q) {
                const float v = __ldg(alts + a * pa.n + q);
#pragma unroll
                for (int i = 0; i < D0; ++i)
                    if (i == pa.col[q]) { same = same && (v == x0[i]); x[i] = v; }
            }
            if (same) continue;
            const bool p1 = fb_pos(fb_spec_forward<S, 0, float>(prm, x, vis));
            if (p1 != p0) flip = a;
        }
        orig_pred[r] = p0 ? 1 : 0;
        first_flip[r] = flip;
    }
}

// ---------------------------------------------------------------------------------------------------
// per-partition sampled counterexample screen (the property the drivers hand to Z3,
// src/AC/Verify-AC.py:137-156): one warp per (partition, chunk), lanes are samples; the protected
// column is swept over the partition's own range.  Result: packed 64-bit key (sample << 32 | fields of
// fb_screen_sample), atomicMin'ed into ce_key[p] (pre-set to ~0).
//
// Relaxed queries (relaxed/AC/Verify-AC.py:48-51, :128-132 with in_const_diff_adult, utils/verif_utils.py:859-867):
// the two points may also differ in one more attribute RA by at most `thr`.  Only the first point has to lie in the
// partition (in_const_domain_* bounds x' on the protected columns alone, :846-854), so the second point's RA value
// ranges over a - thr .. a + thr around the sample's own value a, unclipped.
// ---------------------------------------------------------------------------------------------------
#define FB_SCREEN_SKIPPED 0xfffffffffffffffeull     // key of a box whose protected range does not fit the key's fields
#define FB_SCREEN_NOSEARCH 0xfffffffffffffffdull    // key pre-set by the caller's skip mask: the kernels leave the box alone
#define FB_SCREEN_MAX_W_STRICT 65536u
#define FB_SCREEN_MAX_W_RELAXED 256u
#define FB_SCREEN_MAX_THR 127

// One sample.  ev(v, r) = logit with the protected column set to v and, in relaxed mode, the RA column to r.
// Strict pairs first: the smallest protected value with a positive logit and the smallest with a negative one (low
// word = off_pos << 16 | off_neg, offsets from pa_lo).  Relaxed mode, when the strict sweep shows one class only: the
// first (delta ascending from -thr, protected value ascending) second point of the other class that has a partner of
// a different protected value in the strict sweep (low word = off_pos << 24 | off_neg << 16 | (dpos+128) << 8 | (dneg+128),
// d* = RA offset of that point, 0 for the one inside the partition).  0xffffffff: no counterexample in this sample.
// One sample's sweep, row by row: rows 0 .. W-1 are the strict sweep of the protected column (RA offset 0), then, in
// relaxed mode, W rows per RA offset d = -thr .. thr without 0.  row() is fed the logits in that order; `low` ends as
// the key described above fb_screen_sample, `done` says no later row matters.
struct FbScreenScan {
    int vp, vp2, vn, vn2;                    // first / second protected offset with a positive / negative logit, -1 = none
    uint32_t low;                            // 0xffffffff: no counterexample (yet)
    bool done;
    __device__ __forceinline__ void begin() { vp = vp2 = vn = vn2 = -1; low = 0xffffffffu; done = false; }
    // k = 0: strict sweep; k >= 1: the k-th relaxed offset d
    __device__ __forceinline__ void row(int k, int off, int d, float y, int W, bool relaxed)
    {
        if (done) return;
        if (k == 0) {
            if (y > 0.0f) { if (vp < 0) vp = off; else if (vp2 < 0) vp2 = off; }
            if (y < 0.0f) { if (vn < 0) vn = off; else if (vn2 < 0) vn2 = off; }
            if (off == W - 1) {
                if (vp >= 0 && vn >= 0) {
                    low = relaxed ? ((uint32_t)vp << 24) | ((uint32_t)vn << 16) | (128u << 8) | 128u
                                  : ((uint32_t)vp << 16) | (uint32_t)vn;
                    done = true;
                } else if (!relaxed || (vp < 0) == (vn < 0)) {
                    done = true;                                        // one sweep only, or no class at all
                }
            }
            return;
        }
        if (vp >= 0 && y < 0.0f) {
            const int partner = (vp != off) ? vp : vp2;
            if (partner >= 0) { low = ((uint32_t)partner << 24) | ((uint32_t)off << 16) | (128u << 8) | (uint32_t)(d + 128); done = true; }
        } else if (vn >= 0 && y > 0.0f) {
            const int partner = (vn != off) ? vn : vn2;
            if (partner >= 0) { low = ((uint32_t)off << 24) | ((uint32_t)partner << 16) | ((uint32_t)(d + 128) << 8) | 128u; done = true; }
        }
    }
    // row q of the sweep, q = k * W + off
    __device__ __forceinline__ void row(int q, float y, int W, bool relaxed, int thr)
    {
        const int k = q / W, off = q - k * W;
        int d = k == 0 ? 0 : -thr + (k - 1);
        if (k != 0 && d >= 0) ++d;
        row(k, off, d, y, W, relaxed);
    }
};

template <class Eval>
__device__ __forceinline__ uint32_t fb_screen_sample(Eval &ev, int pa_lo, uint32_t W, bool relaxed, int ra_val, int thr)
{
    // (one loop over all rows, so that the network is inlined once)
    FbScreenScan sc;
    sc.begin();
    const int nk = relaxed ? 2 * thr + 1 : 1;
    int d = 0;
    for (int k = 0; k < nk && !sc.done; ++k) {
        for (int off = 0; off < (int)W && !sc.done; ++off)
            sc.row(k, off, d, ev(pa_lo + off, ra_val + d), (int)W, relaxed);
        d = (k == 0) ? -thr : (d + 1 == 0 ? 1 : d + 1);
    }
    return sc.low;
}

template <class S, bool PHILOX, bool RELAXED>
__global__ void __launch_bounds__(FB_PROFILE_THREADS)
fb_partition_screen_spec_kernel(const __grid_constant__ SpecParams<S, float> prm, int64_t n,
                                const int32_t *__restrict__ boxes, const int32_t *__restrict__ samples, uint64_t seed,
                                int64_t first_index, int nsamp, int chunks, int chunk_len, int pa_col, int ra_col, int ra_thr,
                                unsigned long long *ce_key)
{
    constexpr int D0 = S::D0;
    const int lane = threadIdx.x & 31;
    const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
    const int64_t gwarp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int64_t units = n * chunks;
    constexpr bool relaxed = RELAXED;                    // (the launcher picks the instantiation from ra_col >= 0)
    for (int64_t u = gwarp; u < units; u += nwarps) {
        const int64_t p = u / chunks;
        if (ce_key[p] == FB_SCREEN_NOSEARCH) continue;
        const int c = (int)(u - p * chunks);
        const int s0 = c * chunk_len;
        const int s1 = min(nsamp, s0 + chunk_len);
        int32_t lo[D0];
        uint32_t range[D0];
        int pa_lo = 0;
        uint32_t W = 0;
        {
            const int2 *bx = reinterpret_cast<const int2 *>(boxes + p * (2 * D0));
#pragma unroll
            for (int i = 0; i < D0; ++i) {
                const int2 b = __ldg(bx + i);
                lo[i] = b.x; range[i] = (uint32_t)(b.y - b.x + 1);
                if (i == pa_col) { pa_lo = b.x; W = range[i]; }
            }
        }
        if (W == 0 || W > (relaxed ? FB_SCREEN_MAX_W_RELAXED : FB_SCREEN_MAX_W_STRICT)) {    // (W == 0: a 2^32-wide range)
            if (lane == 0 && c == 0) ce_key[p] = FB_SCREEN_SKIPPED;
            continue;
        }
        uint32_t widest = 0;
#pragma unroll
        for (int i = 0; i < D0; ++i) widest = max(widest, range[i]);
        const int grp = fb_draw_group(widest);
        const int64_t pidx = first_index + p;
        unsigned long long best = ~0ull;
        // sample s of this partition as float32 features (s_src: the row to read when samples are injected)
        auto load = [&](int s, int s_src, float (&x0)[D0]) {
            if (PHILOX) {
                if (grp == 4) fb_draw_spec<D0, 1, 4>(lo, range, s, pidx, seed, x0);
                else if (grp == 2) fb_draw_spec<D0, 1, 2>(lo, range, s, pidx, seed, x0);
                else fb_draw_spec<D0, 1, 1>(lo, range, s, pidx, seed, x0);
            } else {
                const int32_t *row = samples + ((int64_t)p * nsamp + s_src) * D0;
#pragma unroll
                for (int i = 0; i < D0; ++i) x0[i] = (float)__ldg(row + i);
            }
        };
        // Relaxed queries sweep W * (2 thr + 1) points per sample, and in a box without a pair every sample is swept to
        // the end.  So each sample first goes through the network ONCE as an interval: protected column over the box's
        // range, relaxed column over a - thr .. a + thr, the other features at the sample's values (fb_spec_forward on
        // FbIv: the same multiply-adds in the same order on both ends, which encloses every logit the sweep would
        // compute, rounding included).  A one-signed enclosure means the sweep finds one class only: no pair in this
        // sample.  The others wait in `pend` (ascending) until 32 of them fill a sweep step.
        __shared__ int s_pend[FB_PROFILE_THREADS / 32][64];
        int *pend = s_pend[threadIdx.x >> 5];
        int npend = 0, sb = s0;
        // The same enclosure once for the WHOLE box (every feature over its range, the relaxed column widened by thr)
        // before any sample is drawn: one-signed means no sample of this box can show two classes.  Only with generated
        // samples (injected ones are the caller's and need not lie in the box), only for the smaller architectures (the
        // strict kernels of the big ones would otherwise carry an interval pass just for this), and not for boxes whose
        // ranges do not fit the arithmetic below.
        bool boxdone = !(PHILOX && S::P <= FB_SCREEN_SPEC_RELAXED_MAXP);
#pragma unroll
        for (int i = 0; i < D0; ++i)
            if (range[i] == 0u || lo[i] < -(1 << 30) || lo[i] > (1 << 30) || range[i] > (1u << 30)) boxdone = true;
        while (sb < s1 || npend > 0) {
            if ((relaxed && sb < s1 && npend < 32) || !boxdone) {
                const bool boxpass = !boxdone;
                const int s = sb + lane;
                const bool active = s < s1;
                float x0[D0];
                load(s, active ? s : s0, x0);
                FbIv xi[D0];
#pragma unroll
                for (int i = 0; i < D0; ++i) {
                    xi[i].lo = xi[i].hi = x0[i];
                    if (i == pa_col) { xi[i].lo = (float)pa_lo; xi[i].hi = (float)(pa_lo + (int)W - 1); }
                    if (i == ra_col) { const int a = (int)x0[i]; xi[i].lo = (float)(a - ra_thr); xi[i].hi = (float)(a + ra_thr); }
                    if (boxpass) {
                        const int w = (relaxed && i == ra_col) ? ra_thr : 0;
                        xi[i].lo = (float)(lo[i] - w); xi[i].hi = (float)(lo[i] + (int)(range[i] - 1u) + w);
                    }
                }
                FbIvReluVisit vis;
                const FbIv y = fb_spec_forward<S, 0, FbIv>(prm, xi, vis);
                if (boxpass) {
                    boxdone = true;
                    if (y.lo > 0.0f || y.hi < 0.0f) break;           // (the same for every lane: the box, not a sample)
                    continue;
                }
                const bool open = active && !(y.lo > 0.0f || y.hi < 0.0f);
                const uint32_t m = __ballot_sync(0xffffffffu, open);
                if (open) pend[npend + __popc(m & ((1u << lane) - 1u))] = s;
                npend += __popc(m);
                sb += 32;
                __syncwarp();
                continue;
            }
            int s;
            bool active;
            if (relaxed) { active = lane < npend; s = active ? pend[lane] : s0; }
            else { s = sb + lane; active = s < s1; sb += 32; }
            float x0[D0];
            load(s, active ? s : s0, x0);
            int ra_val = 0;
#pragma unroll
            for (int i = 0; i < D0; ++i)
                if (i == ra_col) ra_val = (int)x0[i];
            auto ev = [&](int v, int r) {
                float x[D0];
#pragma unroll
                for (int i = 0; i < D0; ++i) x[i] = (i == pa_col) ? (float)v : ((i == ra_col) ? (float)r : x0[i]);
                FbReluVisit<float> vis;
                return fb_spec_forward<S, 0, float>(prm, x, vis);
            };
            const uint32_t low = fb_screen_sample(ev, pa_lo, W, relaxed, ra_val, ra_thr);
            const bool ce = active && low != 0xffffffffu;
            if (ce && best == ~0ull) best = ((unsigned long long)(uint32_t)s << 32) | low;
            if (__any_sync(0xffffffffu, ce)) break;          // later samples cannot beat this step's
            if (relaxed) {
                const int rest = npend > 32 ? npend - 32 : 0;
                const int v = lane < rest ? pend[32 + lane] : 0;
                __syncwarp();
                if (lane < rest) pend[lane] = v;
                __syncwarp();
                npend = rest;
            }
        }
        // lowest sample index wins; keys are unique per sample so min over the packed key is enough
        const uint32_t hi = (uint32_t)(best >> 32);
        const uint32_t hmin = __reduce_min_sync(0xffffffffu, hi);
        if (hmin != 0xffffffffu && hi == hmin) atomicMin(&ce_key[p], best);
    }
}
