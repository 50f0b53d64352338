// fb_spec_ibp.cuh -- compile-time-specialised interval recurrence + mask algebra (see fb_common.cuh): per box
// 8P + 4N directed/nearest DMUL/DADD, weights as immediate constant-bank operands, bounds in registers.
#pragma once
#include "fb_common.cuh"

// ---------------------------------------------------------------------------------------------------
// interval recurrence + mask algebra, one partition per thread.
// reference: neuron_bounds utils/prune.py:105-164 (round-to-nearest chain, separately rounded product
// and sum), dead_node_from_bound :226-251, keep-one fix :689-691, merge :941-948.  The second chain
// rounds outward (lower bound toward -inf, upper toward +inf) and is what the verdict and the reported
// enclosure use.
// ---------------------------------------------------------------------------------------------------
struct FbPruneOut {
    uint32_t *cand;       // [n][NW]  count == 0
    uint32_t *b_dead;     // [n][NW]  cand & (ws_ub <= 0), keep-one applied
    uint32_t *dead;       // [n][NW]  b_dead | s_dead, keep-one applied
    uint32_t *sliver;     // [n][NW]  neurons whose s_dead is within rounding error: the host Z3 query decides
    uint8_t *verdict;     // [n]
    uint16_t *counts16;   // [n][N]   non-zero activation counts (pos_prob * S)
    double *ws_lb, *ws_ub, *pl_lb, *pl_ub;   // [n][N]  round-to-nearest, reference-exact
    double *lb_rd, *ub_ru;                   // [n][N]  outward rounded
    uint32_t *s_dead;     // [n][NW]  singular verification in closed form (fb_singular below), keep-one applied
    uint32_t *s_cand;     // [n][NW]  b_cand minus the neurons s_dead found (before its keep-one fix)
    uint32_t *record;     // [n][2*NW+2]  verdict | dead[NW] | cand[NW] | first-counterexample sample (-1 here)
};

// singular_verification* (utils/prune.py:276-644) in closed form.  For a still-candidate neuron j of hidden layer
// l >= 1 the reference asks Z3 whether  y_j = sum_k w_kj x_k + b_j > 0  can hold with x_k boxed by
// pl_lb[l][k] <= x_k <= pl_ub[l][k] for k < out_l -- intermediate_domain_constraint (:266-273) is called with the
// layer's OWN index (:302-345), so the bounds are the post-activation bounds of layer l itself, index for index, and the
// inputs k >= out_l are unconstrained Reals.  The query is sat iff an unconstrained input has a non-zero weight (`unb`,
// a per-model bit computed by fb_model_create) or
//     q_j = sum_{k < min(in_l, out_l)} max(w_kj pl_lb[l][k], w_kj pl_ub[l][k]) + b_j > 0,
// so s_dead = b_cand & !unb & (q <= 0).  q is evaluated k ascending with separately rounded product and sum, bias last
// (same as the oracle's singular_one); |q| <= tol flags the neuron for the host's exact Z3 query.  On layer 0 the
// variables are the integer box itself and q is ws_ub, which is > 0 for every b_cand neuron: only the flag remains.
template <class S>
struct FbIbpSink {
    static constexpr int N = S::N, NW = S::NW;
    const uint32_t *cnt;      // this partition's counts or nullptr (then every neuron is a candidate)
    const double *tol;        // [N] or nullptr
    const uint32_t *unb;      // [NW] per-model: neuron has a non-zero weight on an input the query leaves unbounded
    const FbPruneOut &o;
    int64_t p;
    uint32_t cand_w[NW], bdead_w[NW], sliver_w[NW], bcand_w[NW], sdead_w[NW], scand_w[NW], dead_w[NW];
    double out_lb_rd, out_ub_ru;

    __device__ __forceinline__ FbIbpSink(const FbPruneOut &o_, int64_t p_, const uint32_t *cnt_, const double *tol_,
                                         const uint32_t *unb_)
        : cnt(cnt_), tol(tol_), unb(unb_), o(o_), p(p_)
    {
#pragma unroll
        for (int k = 0; k < NW; ++k) {
            cand_w[k] = 0; bdead_w[k] = 0; sliver_w[k] = 0; bcand_w[k] = 0; sdead_w[k] = 0; scand_w[k] = 0; dead_w[k] = 0;
        }
    }

    // idx / last / first are compile-time constants after unrolling (the call sits in fully unrolled loops)
    __device__ __forceinline__ void neuron(int idx, bool last, bool first, double lb, double ub, double lbd, double ubd)
    {
        uint32_t c32 = 0;
        if (cnt) c32 = cnt[idx];
        const bool c = c32 == 0;
        const bool le0 = ub <= 0.0;
        const uint32_t bit = 1u << (idx & 31);
        if (c) cand_w[idx >> 5] |= bit;
        if (last) { if (c) { bdead_w[idx >> 5] |= bit; bcand_w[idx >> 5] |= bit; } out_lb_rd = lbd; out_ub_ru = ubd; }
        else {
            if (c && le0) bdead_w[idx >> 5] |= bit;
            if (c && !le0) bcand_w[idx >> 5] |= bit;
            if (first && tol != nullptr && c && !le0 && ub <= tol[idx]) sliver_w[idx >> 5] |= bit;
        }
        if (o.counts16) o.counts16[p * N + idx] = (uint16_t)c32;
        if (o.ws_lb) o.ws_lb[p * N + idx] = lb;
        if (o.ws_ub) o.ws_ub[p * N + idx] = ub;
        if (o.pl_lb) o.pl_lb[p * N + idx] = lb < 0.0 ? 0.0 : lb;
        if (o.pl_ub) o.pl_ub[p * N + idx] = ub < 0.0 ? 0.0 : ub;
        if (o.lb_rd) o.lb_rd[p * N + idx] = lbd;
        if (o.ub_ru) o.ub_ru[p * N + idx] = ubd;
    }

    __device__ __forceinline__ bool is_bcand(int idx) const { return (bcand_w[idx >> 5] >> (idx & 31)) & 1u; }

    // the closed-form query result of a b_cand neuron of a hidden layer l >= 1
    __device__ __forceinline__ void singular(int idx, double q)
    {
        const uint32_t bit = 1u << (idx & 31);
        if ((__ldg(unb + (idx >> 5)) & bit) != 0u) return;                  // sat whatever the bounds
        if (q <= 0.0) sdead_w[idx >> 5] |= bit;
        if (tol != nullptr && fabs(q) <= tol[idx]) sliver_w[idx >> 5] |= bit;
    }

    // all bits [BO, BO+OUT) of w set?
    template <int BO, int OUT>
    __device__ __forceinline__ bool all_set(const uint32_t (&w)[NW]) const
    {
        bool all1 = true;
#pragma unroll
        for (int k = BO >> 5; k <= (BO + OUT - 1) >> 5; ++k) {
            const int b0 = (k * 32 > BO) ? k * 32 : BO;
            const int b1 = ((k + 1) * 32 < BO + OUT) ? (k + 1) * 32 : BO + OUT;
            const uint32_t m = (b1 - b0 == 32) ? 0xffffffffu : (((1u << (b1 - b0)) - 1u) << (b0 & 31));
            all1 = all1 && ((w[k] & m) == m);
        }
        return all1;
    }

    // end of the layer occupying bits [BO, BO+OUT): s_cand, then the keep-one fixes (utils/prune.py:689-691, :698-700,
    // :703-705: a mask with no 0 in a layer gets its first entry cleared) on b_dead, s_dead and their union
    template <int BO, int OUT, bool LAST>
    __device__ __forceinline__ void layer_done()
    {
        constexpr uint32_t first = 1u << (BO & 31);
#pragma unroll
        for (int k = BO >> 5; k <= (BO + OUT - 1) >> 5; ++k) {
            const int b0 = (k * 32 > BO) ? k * 32 : BO;
            const int b1 = ((k + 1) * 32 < BO + OUT) ? (k + 1) * 32 : BO + OUT;
            const uint32_t m = (b1 - b0 == 32) ? 0xffffffffu : (((1u << (b1 - b0)) - 1u) << (b0 & 31));
            scand_w[k] |= bcand_w[k] & ~sdead_w[k] & m;
        }
        if (LAST) sdead_w[BO >> 5] |= bcand_w[BO >> 5] & first;               // copied through (:287-288), cleared below
        if (all_set<BO, OUT>(bdead_w)) bdead_w[BO >> 5] &= ~first;
        if (all_set<BO, OUT>(sdead_w)) sdead_w[BO >> 5] &= ~first;
#pragma unroll
        for (int k = BO >> 5; k <= (BO + OUT - 1) >> 5; ++k) {
            const int b0 = (k * 32 > BO) ? k * 32 : BO;
            const int b1 = ((k + 1) * 32 < BO + OUT) ? (k + 1) * 32 : BO + OUT;
            const uint32_t m = (b1 - b0 == 32) ? 0xffffffffu : (((1u << (b1 - b0)) - 1u) << (b0 & 31));
            dead_w[k] |= (bdead_w[k] | sdead_w[k]) & m;
        }
        if (all_set<BO, OUT>(dead_w)) dead_w[BO >> 5] &= ~first;
    }

    __device__ __forceinline__ void finish()
    {
        uint32_t anysl = 0;
#pragma unroll
        for (int k = 0; k < NW; ++k) {
            if (o.cand) o.cand[p * NW + k] = cand_w[k];
            if (o.b_dead) o.b_dead[p * NW + k] = bdead_w[k];
            if (o.dead) o.dead[p * NW + k] = dead_w[k];
            if (o.sliver) o.sliver[p * NW + k] = sliver_w[k];
            if (o.s_dead) o.s_dead[p * NW + k] = sdead_w[k];
            if (o.s_cand) o.s_cand[p * NW + k] = scand_w[k];
            anysl |= sliver_w[k];
        }
        uint8_t v = 0;
        if (out_lb_rd > 0.0 || out_ub_ru < 0.0) v |= FB_VERDICT_CERTIFIED;
        if (anysl) v |= FB_VERDICT_SLIVER;
        if (o.verdict) o.verdict[p] = v;
        if (o.record) {
            uint32_t *r = o.record + p * (2 * NW + 2);
            r[0] = v;
#pragma unroll
            for (int k = 0; k < NW; ++k) { r[1 + k] = dead_w[k]; r[1 + NW + k] = cand_w[k]; }
            r[1 + 2 * NW] = 0xffffffffu;
        }
    }
};

template <class S, int l>
__device__ __forceinline__ void fb_spec_ibp(const SpecParams<S, double> &prm, const double (&lb)[S::dim(l)],
                                            const double (&ub)[S::dim(l)], const double (&lbd)[S::dim(l)],
                                            const double (&ubd)[S::dim(l)], FbIbpSink<S> &sink)
{
    constexpr int IN = S::dim(l), OUT = S::dim(l + 1), BO = S::boff(l);
    constexpr bool last = (l == S::NL - 1);
    double nlb[OUT], nub[OUT], nlbd[OUT], nubd[OUT];
#pragma unroll
    for (int j = 0; j < OUT; ++j) {
        double mn = 0.0, mx = 0.0, mnd = 0.0, mxd = 0.0;
#pragma unroll
        for (int i = 0; i < IN; ++i) {
            const double w = prm.w[S::widx(l, i, j)];
            const bool neg = fb_signbit(w);            // host canonicalises -0.0 to +0.0
            const double a = neg ? ub[i] : lb[i], b = neg ? lb[i] : ub[i];
            const double ad = neg ? ubd[i] : lbd[i], bd = neg ? lbd[i] : ubd[i];
            mn = __dadd_rn(mn, __dmul_rn(w, a));
            mx = __dadd_rn(mx, __dmul_rn(w, b));
            mnd = __dadd_rd(mnd, __dmul_rd(w, ad));
            mxd = __dadd_ru(mxd, __dmul_ru(w, bd));
        }
        const double bj = prm.b[BO + j];
        const double wl = __dadd_rn(mn, bj), wu = __dadd_rn(mx, bj);
        const double wld = __dadd_rd(mnd, bj), wud = __dadd_ru(mxd, bj);
        sink.neuron(BO + j, last, l == 0, wl, wu, wld, wud);
        nlb[j] = wl < 0.0 ? 0.0 : wl; nub[j] = wu < 0.0 ? 0.0 : wu;
        nlbd[j] = wld < 0.0 ? 0.0 : wld; nubd[j] = wud < 0.0 ? 0.0 : wud;
    }
    if constexpr (l >= 1 && !last) {
        // singular verification of this layer's remaining candidates (rare: most boxes have none left here)
        constexpr int MM = IN < OUT ? IN : OUT;
        uint32_t any = 0;
#pragma unroll
        for (int k = BO >> 5; k <= (BO + OUT - 1) >> 5; ++k) any |= sink.bcand_w[k];     // (earlier layers' bits are 0 here
        if (any) {                                                                       //  only if they had no candidate: harmless)
#pragma unroll
            for (int j = 0; j < OUT; ++j) {
                if (sink.is_bcand(BO + j)) {
                    double q = 0.0;
#pragma unroll
                    for (int k = 0; k < MM; ++k) {
                        const double w = prm.w[S::widx(l, k, j)];
                        q = __dadd_rn(q, __dmul_rn(w, fb_signbit(w) ? nlb[k] : nub[k]));
                    }
                    sink.singular(BO + j, __dadd_rn(q, (double)prm.b[BO + j]));
                }
            }
        }
    }
    sink.template layer_done<BO, OUT, last>();
    if constexpr (!last) fb_spec_ibp<S, l + 1>(prm, nlb, nub, nlbd, nubd, sink);
}

#define FB_IBP_THREADS 128
#ifndef FB_IBP_MINB
#define FB_IBP_MINB 1
#endif

template <class S>
__global__ void __launch_bounds__(FB_IBP_THREADS, FB_IBP_MINB)
fb_ibp_spec_kernel(const __grid_constant__ SpecParams<S, double> prm, int64_t n, const int32_t *__restrict__ boxes,
                   const uint32_t *__restrict__ counts, const double *__restrict__ tol,
                   const uint32_t *__restrict__ unb, const __grid_constant__ FbPruneOut out)
{
    constexpr int D0 = S::D0;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += stride) {
        double lb[D0], ub[D0];
        const int2 *bx = reinterpret_cast<const int2 *>(boxes + p * (2 * D0));
#pragma unroll
        for (int i = 0; i < D0; ++i) {
            const int2 b = __ldg(bx + i);
            lb[i] = (double)b.x; ub[i] = (double)b.y;
        }
        FbIbpSink<S> sink(out, p, counts ? counts + p * S::N : nullptr, tol, unb);
        fb_spec_ibp<S, 0>(prm, lb, ub, lb, ub, sink);
        sink.finish();
    }
}


// ---------------------------------------------------------------------------------------------------
// The same recurrence with the neuron loop left as a LOOP.
//
// fb_ibp_spec_kernel above is one straight line of ~23 instructions per weight (AC-1: 8 000 instructions, 127 KB):
// it does not fit the instruction cache, and with four bound arrays of two layers live it needs more than 255
// registers (ncu on B200, AC-1: "no instruction" is the top stall, 1 warp per scheduler, 22 % of the FP64 pipe).
// Here only the loop over a neuron's INPUTS is unrolled -- the inputs' bounds sit in registers under compile-time
// indices -- while the neurons of a layer are visited by a run-time loop: the weight is fetched from the constant
// bank at a run-time offset, the finished neuron's post-activation bounds go to this thread's column of a small
// shared-memory array, and the next layer reads them back into registers.  Code is ~20 instructions per INPUT of a
// layer (AC-1: under 1 000 in all), live state 4 x (widest input) doubles.  Masks are kept per layer as 64-bit words
// (bit j = neuron j), which makes the keep-one fixes single comparisons, and merged into the output words at the end
// of the layer.  Same arithmetic, same outputs as the unrolled kernel.  For shapes whose hidden layers are at most
// FB_IBP_LOOP_MAXW wide.
// ---------------------------------------------------------------------------------------------------
#define FB_IBP_LOOP_MAXW 16
#define FB_IBP_LOOP_THREADS 128

template <class S> __host__ __device__ constexpr int fb_ibp_max_hidden()
{
    int r = 0;
    for (int l = 1; l < S::NL; ++l) r = S::dim(l) > r ? S::dim(l) : r;      // inputs of layers 1 .. (hidden widths)
    return r;
}
// networks with ONE hidden layer (any width up to 128) take the fused form below: no hidden bounds are stored at all
template <class S> __host__ __device__ constexpr bool fb_ibp_fused2_ok() { return S::NL == 2 && S::dim(1) <= 128; }
template <class S> __host__ __device__ constexpr bool fb_ibp_loop_ok()
{
    return fb_ibp_fused2_ok<S>() || (fb_ibp_max_hidden<S>() <= FB_IBP_LOOP_MAXW && S::NL >= 2);
}

template <int NW>
struct FbIbpMasks {
    uint32_t cand[NW], bdead[NW], sliver[NW], sdead[NW], scand[NW], dead[NW];
    // bits [BO, BO + 64) of w |= m
    template <int BO>
    static __device__ __forceinline__ void put(uint32_t (&w)[NW], uint64_t m)
    {
        constexpr int k0 = BO >> 5, s = BO & 31;
        const uint64_t lo = m << s;
        w[k0] |= (uint32_t)lo;
        if (k0 + 1 < NW) w[k0 + 1] |= (uint32_t)(lo >> 32);
        if (s != 0 && k0 + 2 < NW) w[k0 + 2] |= (uint32_t)(m >> (64 - s));
    }
};

// FMA0: the box is small enough (|x| < 2^29) that a float32-valued weight times an input of layer 0 has at most 53
// significant bits.  The product is then exact, "round the product, then round the sum" equals one fused multiply-add in
// every rounding mode, and layer 0 -- most of the weights of the small networks -- costs half the FP64 instructions.
template <class S, int l, int INW, bool FMA0>
__device__ __forceinline__ void fb_ibp_loop_layer(const SpecParams<S, double> &prm, const double (&lb)[INW], const double (&ub)[INW],
                                                  const double (&lbd)[INW], const double (&ubd)[INW], double *scol,
                                                  FbIbpMasks<S::NW> &mk, const uint32_t *cnt, const double *tol,
                                                  const uint32_t *unb, const FbPruneOut &o, int64_t p, double &out_lb_rd,
                                                  double &out_ub_ru)
{
    constexpr int IN = S::dim(l), OUT = S::dim(l + 1), BO = S::boff(l), N = S::N;
    constexpr bool last = (l == S::NL - 1), first = (l == 0);
    static_assert(IN <= INW && OUT <= 64, "layer too wide for the looped bounds kernel");
    constexpr uint64_t full = OUT == 64 ? ~0ull : ((1ull << OUT) - 1ull);
    uint64_t cand64 = 0, bdead64 = 0, bcand64 = 0, sliver64 = 0, sdead64 = 0;
    // weights of neuron j: even layers are stored [out][in], odd layers [in][out] (Spec::widx)
    constexpr int jstride = (l & 1) ? 1 : IN, istride = (l & 1) ? OUT : 1;
    const double *wl = prm.w + S::woff(l);
#pragma unroll 1
    for (int j = 0; j < OUT; ++j) {
        const double *wj = wl + j * jstride;
        double mn = 0.0, mx = 0.0, mnd = 0.0, mxd = 0.0;
#pragma unroll
        for (int i = 0; i < IN; ++i) {
            const double w = wj[i * istride];
            const bool neg = fb_signbit(w);            // host canonicalises -0.0 to +0.0
            const double a = neg ? ub[i] : lb[i], b = neg ? lb[i] : ub[i];
            if (first && FMA0) {                       // exact products (see FMA0)
                mn = __fma_rn(w, a, mn); mx = __fma_rn(w, b, mx);
                mnd = __fma_rd(w, a, mnd); mxd = __fma_ru(w, b, mxd);
                continue;
            }
            mn = __dadd_rn(mn, __dmul_rn(w, a));
            mx = __dadd_rn(mx, __dmul_rn(w, b));
            if (first) {                               // integer inputs: both chains read the same box
                mnd = __dadd_rd(mnd, __dmul_rd(w, a));
                mxd = __dadd_ru(mxd, __dmul_ru(w, b));
            } else {
                const double ad = neg ? ubd[i] : lbd[i], bd = neg ? lbd[i] : ubd[i];
                mnd = __dadd_rd(mnd, __dmul_rd(w, ad));
                mxd = __dadd_ru(mxd, __dmul_ru(w, bd));
            }
        }
        const double bj = prm.b[BO + j];
        const double wlo = __dadd_rn(mn, bj), wup = __dadd_rn(mx, bj);
        const double wld = __dadd_rd(mnd, bj), wud = __dadd_ru(mxd, bj);
        const int idx = BO + j;
        const uint32_t c32 = cnt ? cnt[idx] : 0u;
        const bool c = c32 == 0, le0 = wup <= 0.0;
        const uint64_t bit = 1ull << j;
        if (c) cand64 |= bit;
        if (last) { if (c) { bdead64 |= bit; bcand64 |= bit; } out_lb_rd = wld; out_ub_ru = wud; }
        else {
            if (c && le0) bdead64 |= bit;
            if (c && !le0) bcand64 |= bit;
            if (first && tol != nullptr && c && !le0 && wup <= tol[idx]) sliver64 |= bit;
        }
        if (o.counts16) o.counts16[p * N + idx] = (uint16_t)c32;
        if (o.ws_lb) o.ws_lb[p * N + idx] = wlo;
        if (o.ws_ub) o.ws_ub[p * N + idx] = wup;
        if (o.pl_lb) o.pl_lb[p * N + idx] = wlo < 0.0 ? 0.0 : wlo;
        if (o.pl_ub) o.pl_ub[p * N + idx] = wup < 0.0 ? 0.0 : wup;
        if (o.lb_rd) o.lb_rd[p * N + idx] = wld;
        if (o.ub_ru) o.ub_ru[p * N + idx] = wud;
        if (!last) {
            // post-activation bounds of neuron j -> this thread's column: [j][4][threads]
            double *q = scol + j * (4 * FB_IBP_LOOP_THREADS);
            q[0] = wlo < 0.0 ? 0.0 : wlo; q[FB_IBP_LOOP_THREADS] = wup < 0.0 ? 0.0 : wup;
            q[2 * FB_IBP_LOOP_THREADS] = wld < 0.0 ? 0.0 : wld; q[3 * FB_IBP_LOOP_THREADS] = wud < 0.0 ? 0.0 : wud;
        }
    }
    if constexpr (!last) {
        double nlb[INW], nub[INW], nlbd[INW], nubd[INW];
#pragma unroll
        for (int k = 0; k < OUT; ++k) {
            const double *q = scol + k * (4 * FB_IBP_LOOP_THREADS);
            nlb[k] = q[0]; nub[k] = q[FB_IBP_LOOP_THREADS]; nlbd[k] = q[2 * FB_IBP_LOOP_THREADS]; nubd[k] = q[3 * FB_IBP_LOOP_THREADS];
        }
        if constexpr (l >= 1) {
            // singular verification of this layer's remaining candidates (see FbIbpSink): inputs = the layer's own bounds
            constexpr int MM = IN < OUT ? IN : OUT;
            uint64_t todo = bcand64;
            while (todo) {
                const int j = __ffsll((long long)todo) - 1;
                todo &= todo - 1;
                const int idx = BO + j;
                if ((__ldg(unb + (idx >> 5)) >> (idx & 31)) & 1u) continue;           // sat whatever the bounds
                const double *wj = wl + j * jstride;
                double q = 0.0;
#pragma unroll
                for (int k = 0; k < MM; ++k) {
                    const double w = wj[k * istride];
                    q = __dadd_rn(q, __dmul_rn(w, fb_signbit(w) ? nlb[k] : nub[k]));
                }
                q = __dadd_rn(q, (double)prm.b[idx]);
                if (q <= 0.0) sdead64 |= 1ull << j;
                if (tol != nullptr && fabs(q) <= tol[idx]) sliver64 |= 1ull << j;
            }
        }
        const uint64_t scand64 = bcand64 & ~sdead64;
        if (bdead64 == full) bdead64 &= ~1ull;               // keep-one fixes (utils/prune.py:689-691, :698-700, :703-705)
        if (sdead64 == full) sdead64 &= ~1ull;
        uint64_t dead64 = bdead64 | sdead64;
        if (dead64 == full) dead64 &= ~1ull;
        FbIbpMasks<S::NW>::template put<BO>(mk.cand, cand64); FbIbpMasks<S::NW>::template put<BO>(mk.bdead, bdead64);
        FbIbpMasks<S::NW>::template put<BO>(mk.sliver, sliver64); FbIbpMasks<S::NW>::template put<BO>(mk.sdead, sdead64);
        FbIbpMasks<S::NW>::template put<BO>(mk.scand, scand64); FbIbpMasks<S::NW>::template put<BO>(mk.dead, dead64);
        fb_ibp_loop_layer<S, l + 1, INW, FMA0>(prm, nlb, nub, nlbd, nubd, scol, mk, cnt, tol, unb, o, p, out_lb_rd, out_ub_ru);
    } else {
        // the single output neuron: candidate copied through, then cleared by the keep-one fix of its 1-element list
        FbIbpMasks<S::NW>::template put<BO>(mk.cand, cand64);
        FbIbpMasks<S::NW>::template put<BO>(mk.scand, bcand64);
    }
}

// One hidden layer (d - H - 1): every finished hidden neuron is folded straight into the four accumulators of the
// logit (terms in ascending neuron order, separately rounded product and sum -- the order and the arithmetic of
// neuron_bounds), so the hidden layer's bounds never sit anywhere: live state is the box and four doubles.  Neurons
// [64 * HALF, 64 * HALF + 64) of the hidden layer; returns whether all of them ended up in b_dead (keep-one fix).
template <class S, bool FMA0, int HALF>
__device__ __forceinline__ bool fb_ibp_fused2_half(const SpecParams<S, double> &prm, const double (&lb)[S::D0], const double (&ub)[S::D0],
                                                   FbIbpMasks<S::NW> &mk, const uint32_t *cnt, const double *tol, const FbPruneOut &o,
                                                   int64_t p, double &omn, double &omx, double &omnd, double &omxd)
{
    constexpr int IN = S::D0, H = S::dim(1), N = S::N;
    constexpr int J0 = 64 * HALF, J1 = (H < J0 + 64) ? H : J0 + 64, CNT = J1 - J0;
    constexpr uint64_t full = CNT == 64 ? ~0ull : ((1ull << CNT) - 1ull);
    uint64_t cand64 = 0, bdead64 = 0, bcand64 = 0, sliver64 = 0;
    const double *w0 = prm.w + S::woff(0);              // layer 0 is stored [out][in]
    const double *w1 = prm.w + S::woff(1);              // layer 1 [in][out] with out = 1
#pragma unroll 1
    for (int j = J0; j < J1; ++j) {
        const double *wj = w0 + j * IN;
        double mn = 0.0, mx = 0.0, mnd = 0.0, mxd = 0.0;
#pragma unroll
        for (int i = 0; i < IN; ++i) {
            const double w = wj[i];
            const bool neg = fb_signbit(w);
            const double a = neg ? ub[i] : lb[i], b = neg ? lb[i] : ub[i];
            if (FMA0) {
                mn = __fma_rn(w, a, mn); mx = __fma_rn(w, b, mx); mnd = __fma_rd(w, a, mnd); mxd = __fma_ru(w, b, mxd);
            } else {
                mn = __dadd_rn(mn, __dmul_rn(w, a)); mx = __dadd_rn(mx, __dmul_rn(w, b));
                mnd = __dadd_rd(mnd, __dmul_rd(w, a)); mxd = __dadd_ru(mxd, __dmul_ru(w, b));
            }
        }
        const double bj = prm.b[j];
        const double wlo = __dadd_rn(mn, bj), wup = __dadd_rn(mx, bj), wld = __dadd_rd(mnd, bj), wud = __dadd_ru(mxd, bj);
        const uint32_t c32 = cnt ? cnt[j] : 0u;
        const bool c = c32 == 0, le0 = wup <= 0.0;
        const uint64_t bit = 1ull << (j - J0);
        if (c) cand64 |= bit;
        if (c && le0) bdead64 |= bit;
        if (c && !le0) bcand64 |= bit;
        if (tol != nullptr && c && !le0 && wup <= tol[j]) sliver64 |= bit;
        const double plb = wlo < 0.0 ? 0.0 : wlo, pub = wup < 0.0 ? 0.0 : wup;
        const double plbd = wld < 0.0 ? 0.0 : wld, pubd = wud < 0.0 ? 0.0 : wud;
        if (o.counts16) o.counts16[p * N + j] = (uint16_t)c32;
        if (o.ws_lb) o.ws_lb[p * N + j] = wlo;
        if (o.ws_ub) o.ws_ub[p * N + j] = wup;
        if (o.pl_lb) o.pl_lb[p * N + j] = plb;
        if (o.pl_ub) o.pl_ub[p * N + j] = pub;
        if (o.lb_rd) o.lb_rd[p * N + j] = wld;
        if (o.ub_ru) o.ub_ru[p * N + j] = wud;
        const double v = w1[j];
        const bool vneg = fb_signbit(v);
        omn = __dadd_rn(omn, __dmul_rn(v, vneg ? pub : plb));
        omx = __dadd_rn(omx, __dmul_rn(v, vneg ? plb : pub));
        omnd = __dadd_rd(omnd, __dmul_rd(v, vneg ? pubd : plbd));
        omxd = __dadd_ru(omxd, __dmul_ru(v, vneg ? plbd : pubd));
    }
    FbIbpMasks<S::NW>::template put<J0>(mk.cand, cand64); FbIbpMasks<S::NW>::template put<J0>(mk.bdead, bdead64);
    FbIbpMasks<S::NW>::template put<J0>(mk.sliver, sliver64); FbIbpMasks<S::NW>::template put<J0>(mk.scand, bcand64);
    return bdead64 == full;
}

template <class S, bool FMA0>
__device__ __forceinline__ void fb_ibp_fused2(const SpecParams<S, double> &prm, const double (&lb)[S::D0], const double (&ub)[S::D0],
                                              FbIbpMasks<S::NW> &mk, const uint32_t *cnt, const double *tol, const FbPruneOut &o,
                                              int64_t p, double &out_lb_rd, double &out_ub_ru)
{
    constexpr int H = S::dim(1), N = S::N, NW = S::NW;
    double omn = 0.0, omx = 0.0, omnd = 0.0, omxd = 0.0;
    bool all = fb_ibp_fused2_half<S, FMA0, 0>(prm, lb, ub, mk, cnt, tol, o, p, omn, omx, omnd, omxd);
    if constexpr (H > 64) all = fb_ibp_fused2_half<S, FMA0, 1>(prm, lb, ub, mk, cnt, tol, o, p, omn, omx, omnd, omxd) && all;
    if (all) mk.bdead[0] &= ~1u;                                  // keep-one fix of the hidden layer (bit 0 = neuron 0)
#pragma unroll
    for (int k = 0; k < NW; ++k) mk.dead[k] = mk.bdead[k];       // s_dead is empty on layer 0 (its query's quantity is ws_ub > 0)
    // the logit
    const double bo = prm.b[H];
    const double wlo = __dadd_rn(omn, bo), wup = __dadd_rn(omx, bo), wld = __dadd_rd(omnd, bo), wud = __dadd_ru(omxd, bo);
    const uint32_t c32 = cnt ? cnt[H] : 0u;
    if (c32 == 0) {                                               // candidate copied through; keep-one clears its dead bits
        mk.cand[H >> 5] |= 1u << (H & 31);
        mk.scand[H >> 5] |= 1u << (H & 31);
    }
    out_lb_rd = wld; out_ub_ru = wud;
    if (o.counts16) o.counts16[p * N + H] = (uint16_t)c32;
    if (o.ws_lb) o.ws_lb[p * N + H] = wlo;
    if (o.ws_ub) o.ws_ub[p * N + H] = wup;
    if (o.pl_lb) o.pl_lb[p * N + H] = wlo < 0.0 ? 0.0 : wlo;
    if (o.pl_ub) o.pl_ub[p * N + H] = wup < 0.0 ? 0.0 : wup;
    if (o.lb_rd) o.lb_rd[p * N + H] = wld;
    if (o.ub_ru) o.ub_ru[p * N + H] = wud;
}

template <class S>
__global__ void __launch_bounds__(FB_IBP_LOOP_THREADS)
fb_ibp_loop_kernel(const __grid_constant__ SpecParams<S, double> prm, int64_t n, const int32_t *__restrict__ boxes,
                   const uint32_t *__restrict__ counts, const double *__restrict__ tol,
                   const uint32_t *__restrict__ unb, const __grid_constant__ FbPruneOut out)
{
    constexpr int D0 = S::D0, NW = S::NW, N = S::N;
    constexpr bool FUSED2 = fb_ibp_fused2_ok<S>();
    constexpr int INW = (FUSED2 || fb_ibp_max_hidden<S>() <= D0) ? D0 : fb_ibp_max_hidden<S>();
    extern __shared__ __align__(16) unsigned char fb_smem[];                 // [max hidden width][4][threads] doubles
    double *scol = reinterpret_cast<double *>(fb_smem) + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += stride) {
        double lb[INW], ub[INW];
        const int2 *bx = reinterpret_cast<const int2 *>(boxes + p * (2 * D0));
        bool small = true;
#pragma unroll
        for (int i = 0; i < D0; ++i) {
            const int2 b = __ldg(bx + i);
            lb[i] = (double)b.x; ub[i] = (double)b.y;
            small = small && b.x > -(1 << 29) && b.x < (1 << 29) && b.y > -(1 << 29) && b.y < (1 << 29);
        }
        FbIbpMasks<NW> mk;
#pragma unroll
        for (int k = 0; k < NW; ++k) { mk.cand[k] = 0; mk.bdead[k] = 0; mk.sliver[k] = 0; mk.sdead[k] = 0; mk.scand[k] = 0; mk.dead[k] = 0; }
        double out_lb_rd = 0.0, out_ub_ru = 0.0;
        // (two copies of the network's code, chosen by a real branch: ptxas does not if-convert across the neuron loops)
        if constexpr (FUSED2) {
            if (small) fb_ibp_fused2<S, true>(prm, lb, ub, mk, counts ? counts + p * N : nullptr, tol, out, p, out_lb_rd, out_ub_ru);
            else fb_ibp_fused2<S, false>(prm, lb, ub, mk, counts ? counts + p * N : nullptr, tol, out, p, out_lb_rd, out_ub_ru);
        } else {
            if (small)
                fb_ibp_loop_layer<S, 0, INW, true>(prm, lb, ub, lb, ub, scol, mk, counts ? counts + p * N : nullptr, tol, unb, out, p,
                                                   out_lb_rd, out_ub_ru);
            else
                fb_ibp_loop_layer<S, 0, INW, false>(prm, lb, ub, lb, ub, scol, mk, counts ? counts + p * N : nullptr, tol, unb, out, p,
                                                    out_lb_rd, out_ub_ru);
        }
        uint32_t anysl = 0;
#pragma unroll
        for (int k = 0; k < NW; ++k) {
            if (out.cand) out.cand[p * NW + k] = mk.cand[k];
            if (out.b_dead) out.b_dead[p * NW + k] = mk.bdead[k];
            if (out.dead) out.dead[p * NW + k] = mk.dead[k];
            if (out.sliver) out.sliver[p * NW + k] = mk.sliver[k];
            if (out.s_dead) out.s_dead[p * NW + k] = mk.sdead[k];
            if (out.s_cand) out.s_cand[p * NW + k] = mk.scand[k];
            anysl |= mk.sliver[k];
        }
        uint8_t v = 0;
        if (out_lb_rd > 0.0 || out_ub_ru < 0.0) v |= FB_VERDICT_CERTIFIED;
        if (anysl) v |= FB_VERDICT_SLIVER;
        if (out.verdict) out.verdict[p] = v;
        if (out.record) {
            uint32_t *r = out.record + p * (2 * NW + 2);
            r[0] = v;
#pragma unroll
            for (int k = 0; k < NW; ++k) { r[1 + k] = mk.dead[k]; r[1 + NW + k] = mk.cand[k]; }
            r[1 + 2 * NW] = 0xffffffffu;
        }
    }
}
