// fb_spec.cuh -- compile-time-specialised kernels (see fb_common.cuh).
//
// Every loop over neurons / inputs is fully unrolled; weights are immediate constant-bank operands
// (the parameter struct is __grid_constant__), activations and interval bounds are registers.  The FP64
// pipe is the only unit these kernels lean on: per sample P DFMA + N DADD (profile), per box
// 8P + 4N directed/nearest DMUL/DADD (interval recurrence).
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

// ---------------------------------------------------------------------------------------------------
// Activation profile.  reference: candidate_dead_nodes, utils/prune.py:168-192 (activation != 0 counted per
// neuron, output layer included) over the samples of simluate_data (:205-222).
//
// One warp per (partition, sample chunk).  Each lane carries K samples per step (VecD<K>: 2 where the register
// budget allows, else 1), so a step covers 32*K samples and every weight fetched from the constant bank feeds K
// fused multiply-adds.  The network itself is fb_spec_forward (layer-pair fused, terms in ascending input order).
// ---------------------------------------------------------------------------------------------------
#ifndef FB_PROFILE_THREADS
#define FB_PROFILE_THREADS 512
#endif
#ifndef FB_PROFILE_SPT
#define FB_PROFILE_SPT 2                                 // samples per lane per step
#endif
#ifndef FB_PROFILE_SPT_LIVE_MAX
#define FB_PROFILE_SPT_LIVE_MAX 26                       // inputs + next-layer accumulators per sample that still fit
#endif
#ifndef FB_PROFILE_SMEM_BOX
#define FB_PROFILE_SMEM_BOX 1
#endif
#ifndef FB_PROFILE_SMEM_BOX_MAXN
#define FB_PROFILE_SMEM_BOX_MAXN 64
#endif
#ifndef FB_PROFILE_MINB
#define FB_PROFILE_MINB 1                                // one 512-thread CTA per SM: caps registers at 128
#endif
#define FB_PROFILE_MAX_STEPS (127 / FB_PROFILE_SPT)      // per-lane counters hold up to 127 (7 bit planes)

#define FB_PROFILE_PLANES 7          // bit-sliced counters count to 127 per lane

// HI_SIGN: hidden neurons' "pre-activation > 0" read from the high word (fb_pos_hi; needs the subnormal-free guarantee
// fb_model_create establishes for every model it gives to these kernels)
template <int N, int K, bool HI_SIGN>
struct FbCountVisit {
    // One bit per neuron and sample: neuron idx sets bit idx of nzmask[k].  After the forward pass the masks
    // are added into `planes`, a bit-sliced counter (planes[b] holds bit b of every neuron's count): two
    // logic ops per plane add one sample to ALL neurons' counters at once, and the activity flag of a ragged
    // step is applied to the whole mask in one AND instead of once per neuron.
    static constexpr int NW = (N + 31) / 32;
    uint32_t nzmask[K][NW];
    __device__ __forceinline__ void begin()
    {
#pragma unroll
        for (int k = 0; k < K; ++k)
#pragma unroll
            for (int w = 0; w < NW; ++w) nzmask[k][w] = 0;
    }
    __device__ __forceinline__ VecD<K> neuron(int idx, bool last, const VecD<K> &pre)
    {
        VecD<K> r;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const bool pos = HI_SIGN ? fb_pos_hi(pre.v[k]) : fb_pos(pre.v[k]);
            const bool nz = last ? fb_nonzero(pre.v[k]) : pos;
            if (nz) nzmask[k][idx >> 5] |= 1u << (idx & 31);
            r.v[k] = last ? pre.v[k] : (pos ? pre.v[k] : 0.0);
        }
        return r;
    }
    // planes += nzmask[k] (restricted to active samples)
    __device__ __forceinline__ void commit(uint32_t (&planes)[FB_PROFILE_PLANES][NW], const bool (&active)[K])
    {
#pragma unroll
        for (int k = 0; k < K; ++k)
#pragma unroll
            for (int w = 0; w < NW; ++w) {
                uint32_t carry = active[k] ? nzmask[k][w] : 0u;
#pragma unroll
                for (int b = 0; b < FB_PROFILE_PLANES; ++b) {
                    const uint32_t t = planes[b][w] & carry;
                    planes[b][w] ^= carry;
                    carry = t;
                }
            }
    }
};

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

// one step of the profile: this lane's K samples (indices s, s+32, ...) through the whole network
template <class S, int K, bool PHILOX, int G, class Visit>
__device__ __forceinline__ void fb_profile_step(const SpecParams<S, double> &prm, Visit &vis, const int32_t (&lo)[S::D0],
                                                const uint32_t (&range)[S::D0], const int32_t *part_samples, int s,
                                                int64_t pidx, uint64_t seed, int s_clamp = 0x7fffffff)
{
    constexpr int D0 = S::D0;
    VecD<K> x[D0];
    if (PHILOX) {
        fb_draw_spec<D0, K, G>(lo, range, s, pidx, seed, x);
    } else {
        const int32_t *row[K];
#pragma unroll
        for (int k = 0; k < K; ++k) row[k] = part_samples + (int64_t)min(s + 32 * k, s_clamp) * D0;
#pragma unroll
        for (int i = 0; i < D0; ++i)
#pragma unroll
            for (int k = 0; k < K; ++k) x[i].v[k] = (double)__ldg(row[k] + i);
    }
    fb_spec_forward<S, 0, VecD<K>>(prm, x, vis);
}

// Samples per lane for shape S: FB_PROFILE_SPT (2) when the live state of the widest layer pair -- the inputs of
// layer l plus the accumulators of layer l+1, per sample -- fits the 128-register budget, else 1.
template <class S>
__host__ __device__ constexpr int fb_profile_spt()
{
    int worst = 0;
    for (int l = 0; l < S::NL; l += 2) {
        const int live = S::dim(l) + (l + 2 <= S::NL ? S::dim(l + 2) : 1);
        if (live > worst) worst = live;
    }
    return worst <= FB_PROFILE_SPT_LIVE_MAX ? FB_PROFILE_SPT : 1;
}

// Whether the partition's box waits in shared memory between steps (see the kernel).  Measured on B200: +6 points of
// FP64 utilisation for the shapes with up to ~50 neurons, -9 for the 100-wide single-hidden-layer shapes.
template <class S>
__host__ __device__ constexpr bool fb_profile_smem_box()
{
    return FB_PROFILE_SMEM_BOX && (S::N <= FB_PROFILE_SMEM_BOX_MAXN || FbTune<S>::smem_box_wide);
}

// The one-instruction sign test pays for the small shapes; on the 100-wide and the deep funnel shapes it changes
// ptxas' schedule for the worse (more weights hoisted into uniform registers and spilled: AC-2 63 -> 59 %, AC-7
// 48 -> 43 % of FP64 peak on B200), so those keep the full compare.
template <class S>
__host__ __device__ constexpr bool fb_profile_hi_sign()
{
    return fb_profile_smem_box<S>();
}

// all steps of one (partition, sample chunk): samples [s0, s1), 32*K per step, counted into the bit-sliced planes
template <class S, int K, bool PHILOX, int G, bool SMEM_BOX>
__device__ __forceinline__ void fb_profile_steps(const SpecParams<S, double> &prm,
                                                 uint32_t (&planes)[FB_PROFILE_PLANES][(S::N + 31) / 32],
                                                 const int32_t (&lo)[S::D0], const uint32_t (&range)[S::D0],
                                                 const int32_t *smem_lo, const uint32_t *smem_range,
                                                 const int32_t *part_samples, int s0, int s1, int lane, int64_t pidx,
                                                 uint64_t seed)
{
    for (int sb = s0; sb < s1; sb += 32 * K) {
        FbCountVisit<S::N, K, fb_profile_hi_sign<S>()> vis;
        vis.begin();
        bool active[K];
#pragma unroll
        for (int k = 0; k < K; ++k) active[k] = (sb + 32 * k + lane) < s1;
        // The barrier keeps loop-invariant loads inside the step.  Without it the compiler hoists constant-bank weight
        // loads (and the shared-memory box) out of the sample loop, runs out of uniform registers and spills them.
        asm volatile("" ::: "memory");
        if constexpr (SMEM_BOX && PHILOX) {
            int32_t rlo[S::D0];
            uint32_t rrange[S::D0];
#pragma unroll
            for (int i = 0; i < S::D0; ++i) { rlo[i] = smem_lo[i]; rrange[i] = smem_range[i]; }
            fb_profile_step<S, K, PHILOX, G>(prm, vis, rlo, rrange, part_samples, sb + lane, pidx, seed, s1 - 1);
        } else
            fb_profile_step<S, K, PHILOX, G>(prm, vis, lo, range, part_samples, sb + lane, pidx, seed, s1 - 1);
        vis.commit(planes, active);
    }
}

// counts[p][N] (uint32): stored when chunks == 1, atomically added otherwise (caller zeroes first).
//
// With PHILOX the kernel exists once per draw class G and a launch handles exactly the partitions whose box has that
// class (the launcher runs G = 4, 2, 1 one after the other; every partition is taken by one of them).  Work is handed
// out in blocks of `ublock` (1 ... 32) consecutive units: lane t classifies unit t of the block, so skipping a block
// that belongs to another launch costs one round of loads per 32 partitions.
template <class S, bool PHILOX, int G>
__global__ void __launch_bounds__(FB_PROFILE_THREADS, FB_PROFILE_MINB)
fb_profile_spec_kernel(const __grid_constant__ SpecParams<S, double> prm, int64_t n,
                       const int32_t *__restrict__ boxes, const int32_t *__restrict__ samples, uint64_t seed,
                       int64_t first_index, int nsamp, int chunks, int chunk_len, int ublock,
                       uint32_t *__restrict__ counts)
{
    constexpr int D0 = S::D0, N = S::N, K = fb_profile_spt<S>();
    const int lane = threadIdx.x & 31;
    const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
    const int64_t gwarp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int64_t units = n * chunks;
    const int64_t ublocks = (units + ublock - 1) / ublock;
    int64_t ub = gwarp, ubase = 0;                    // next block to classify; first unit of the block being worked on
    uint32_t todo = 0;                                // units of that block still to do
    for (;;) {
        while (todo == 0) {
            if (ub >= ublocks) return;
            ubase = ub * ublock;
            const int64_t ut = ubase + lane;
            bool mine = lane < ublock && ut < units;
            if (PHILOX && mine) mine = fb_draw_group(fb_box_widest(boxes + (ut / chunks) * (2 * D0), D0)) == G;
            todo = __ballot_sync(0xffffffffu, mine);
            ub += nwarps;
        }
        const int64_t u = ubase + (__ffs(todo) - 1);
        todo &= todo - 1;
        const int64_t p = u / chunks;
        const int c = (int)(u - p * chunks);
        const int s0 = c * chunk_len;
        const int s1 = min(nsamp, s0 + chunk_len);
        // The box (lo, range per feature) is only needed while a step's samples are drawn.  For most shapes it waits
        // in shared memory (one copy per warp) and is re-read at the top of every step, instead of holding 2*D0
        // registers through the network; SMEM_BOX is false for the shapes where that measured slower.
        constexpr bool SMEM_BOX = fb_profile_smem_box<S>();
        constexpr int D0P = (D0 + 3) & ~3;               // rows padded to 16 bytes: read back as 128-bit words
        __shared__ __align__(16) int32_t s_lo[SMEM_BOX ? FB_PROFILE_THREADS / 32 : 1][SMEM_BOX ? D0P : 4];
        __shared__ __align__(16) uint32_t s_range[SMEM_BOX ? FB_PROFILE_THREADS / 32 : 1][SMEM_BOX ? D0P : 4];
        int32_t lo[D0];
        uint32_t range[D0];
        if (PHILOX) {
            const int2 *bx = reinterpret_cast<const int2 *>(boxes + p * (2 * D0));
            if constexpr (SMEM_BOX) {
                __syncwarp();
                for (int i = lane; i < D0; i += 32) {
                    const int2 b = __ldg(bx + i);
                    s_lo[threadIdx.x >> 5][i] = b.x; s_range[threadIdx.x >> 5][i] = (uint32_t)(b.y - b.x + 1);
                }
                __syncwarp();
            } else {
#pragma unroll
                for (int i = 0; i < D0; ++i) {
                    const int2 b = __ldg(bx + i);
                    lo[i] = b.x; range[i] = (uint32_t)(b.y - b.x + 1);
                }
            }
        }
        const int64_t pidx = first_index + p;
        constexpr int NWm = (N + 31) / 32;
        uint32_t planes[FB_PROFILE_PLANES][NWm];
#pragma unroll
        for (int b = 0; b < FB_PROFILE_PLANES; ++b)
#pragma unroll
            for (int w = 0; w < NWm; ++w) planes[b][w] = 0;
        fb_profile_steps<S, K, PHILOX, G, SMEM_BOX>(prm, planes, lo, range, s_lo[SMEM_BOX ? threadIdx.x >> 5 : 0],
                                                    s_range[SMEM_BOX ? threadIdx.x >> 5 : 0],
                                                    samples + (int64_t)p * nsamp * D0, s0, s1, lane, pidx, seed);
        uint32_t held = 0;
#pragma unroll
        for (int j = 0; j < N; ++j) {
            uint32_t mycount = 0;
#pragma unroll
            for (int b = 0; b < FB_PROFILE_PLANES; ++b) mycount |= ((planes[b][j >> 5] >> (j & 31)) & 1u) << b;
            const uint32_t tot = __reduce_add_sync(0xffffffffu, mycount);
            if ((j & 31) == lane) held = tot;
            if ((j & 31) == 31 || j == N - 1) {
                const int jj = (j & ~31) + lane;
                if (jj <= j) {
                    if (chunks == 1) counts[p * N + jj] = held;
                    else atomicAdd(&counts[p * N + jj], held);
                }
            }
        }
    }
}

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
    uint32_t *dead;       // [n][NW]  b_dead | s_dead (s_dead == 0 on device), keep-one applied
    uint32_t *sliver;     // [n][NW]  neurons whose s_dead needs the host Z3 query
    uint8_t *verdict;     // [n]
    uint16_t *counts16;   // [n][N]   non-zero activation counts (pos_prob * S)
    double *ws_lb, *ws_ub, *pl_lb, *pl_ub;   // [n][N]  round-to-nearest, reference-exact
    double *lb_rd, *ub_ru;                   // [n][N]  outward rounded
};

template <class S>
struct FbIbpSink {
    static constexpr int N = S::N, NW = S::NW;
    const uint32_t *cnt;      // this partition's counts or nullptr (then every neuron is a candidate)
    const double *tol;        // [N] or nullptr
    const FbPruneOut &o;
    int64_t p;
    uint32_t cand_w[NW], bdead_w[NW], sliver_w[NW];
    double out_lb_rd, out_ub_ru;

    __device__ __forceinline__ FbIbpSink(const FbPruneOut &o_, int64_t p_, const uint32_t *cnt_, const double *tol_)
        : cnt(cnt_), tol(tol_), o(o_), p(p_)
    {
#pragma unroll
        for (int k = 0; k < NW; ++k) { cand_w[k] = 0; bdead_w[k] = 0; sliver_w[k] = 0; }
    }

    // idx / last are compile-time constants after unrolling (the call sits in fully unrolled loops)
    __device__ __forceinline__ void neuron(int idx, bool last, double lb, double ub, double lbd, double ubd)
    {
        uint32_t c32 = 0;
        if (cnt) c32 = cnt[idx];
        const bool c = c32 == 0;
        const bool le0 = ub <= 0.0;
        const uint32_t bit = 1u << (idx & 31);
        if (c) cand_w[idx >> 5] |= bit;
        if (last) { if (c) bdead_w[idx >> 5] |= bit; out_lb_rd = lbd; out_ub_ru = ubd; }
        else {
            if (c && le0) bdead_w[idx >> 5] |= bit;
            if (tol != nullptr && c && !le0 && ub <= tol[idx]) sliver_w[idx >> 5] |= bit;
        }
        if (o.counts16) o.counts16[p * N + idx] = (uint16_t)c32;
        if (o.ws_lb) o.ws_lb[p * N + idx] = lb;
        if (o.ws_ub) o.ws_ub[p * N + idx] = ub;
        if (o.pl_lb) o.pl_lb[p * N + idx] = lb < 0.0 ? 0.0 : lb;
        if (o.pl_ub) o.pl_ub[p * N + idx] = ub < 0.0 ? 0.0 : ub;
        if (o.lb_rd) o.lb_rd[p * N + idx] = lbd;
        if (o.ub_ru) o.ub_ru[p * N + idx] = ubd;
    }

    // keep-one fix for the layer occupying bits [BO, BO+OUT)
    template <int BO, int OUT>
    __device__ __forceinline__ void keep_one()
    {
        bool all1 = true;
#pragma unroll
        for (int k = BO >> 5; k <= (BO + OUT - 1) >> 5; ++k) {
            const int b0 = (k * 32 > BO) ? k * 32 : BO;
            const int b1 = ((k + 1) * 32 < BO + OUT) ? (k + 1) * 32 : BO + OUT;
            const uint32_t m = (b1 - b0 == 32) ? 0xffffffffu : (((1u << (b1 - b0)) - 1u) << (b0 & 31));
            all1 = all1 && ((bdead_w[k] & m) == m);
        }
        if (all1) bdead_w[BO >> 5] &= ~(1u << (BO & 31));
    }

    __device__ __forceinline__ void finish()
    {
        uint32_t anysl = 0;
#pragma unroll
        for (int k = 0; k < NW; ++k) {
            if (o.cand) o.cand[p * NW + k] = cand_w[k];
            if (o.b_dead) o.b_dead[p * NW + k] = bdead_w[k];
            if (o.dead) o.dead[p * NW + k] = bdead_w[k];
            if (o.sliver) o.sliver[p * NW + k] = sliver_w[k];
            anysl |= sliver_w[k];
        }
        if (o.verdict) {
            uint8_t v = 0;
            if (out_lb_rd > 0.0 || out_ub_ru < 0.0) v |= FB_VERDICT_CERTIFIED;
            if (anysl) v |= FB_VERDICT_SLIVER;
            o.verdict[p] = v;
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
        sink.neuron(BO + j, last, wl, wu, wld, wud);
        nlb[j] = wl < 0.0 ? 0.0 : wl; nub[j] = wu < 0.0 ? 0.0 : wu;
        nlbd[j] = wld < 0.0 ? 0.0 : wld; nubd[j] = wud < 0.0 ? 0.0 : wud;
    }
    sink.template keep_one<BO, OUT>();
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
                   const __grid_constant__ FbPruneOut out)
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
        FbIbpSink<S> sink(out, p, counts ? counts + p * S::N : nullptr, tol);
        fb_spec_ibp<S, 0>(prm, lb, ub, lb, ub, sink);
        sink.finish();
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
// column is swept over the partition's own range.  Result: packed 64-bit key
// (sample << 32 | (v_pos - lo) << 16 | (v_neg - lo)), atomicMin'ed into ce_key[p] (pre-set to ~0).
// ---------------------------------------------------------------------------------------------------
template <class S, bool PHILOX>
__global__ void __launch_bounds__(FB_PROFILE_THREADS)
fb_partition_screen_spec_kernel(const __grid_constant__ SpecParams<S, float> prm, int64_t n,
                                const int32_t *__restrict__ boxes, const int32_t *__restrict__ samples, uint64_t seed,
                                int64_t first_index, int nsamp, int chunks, int chunk_len, int pa_col,
                                unsigned long long *__restrict__ ce_key)
{
    constexpr int D0 = S::D0;
    const int lane = threadIdx.x & 31;
    const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
    const int64_t gwarp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int64_t units = n * chunks;
    for (int64_t u = gwarp; u < units; u += nwarps) {
        const int64_t p = u / chunks;
        const int c = (int)(u - p * chunks);
        const int s0 = c * chunk_len;
        const int s1 = min(nsamp, s0 + chunk_len);
        int32_t lo[D0];
        uint32_t range[D0];
        int pa_lo = 0, pa_hi = 0;
        {
            const int2 *bx = reinterpret_cast<const int2 *>(boxes + p * (2 * D0));
#pragma unroll
            for (int i = 0; i < D0; ++i) {
                const int2 b = __ldg(bx + i);
                lo[i] = b.x; range[i] = (uint32_t)(b.y - b.x + 1);
                if (i == pa_col) { pa_lo = b.x; pa_hi = b.y; }
            }
        }
        uint32_t widest = 0;
#pragma unroll
        for (int i = 0; i < D0; ++i) widest = max(widest, range[i]);
        const int grp = fb_draw_group(widest);
        const int64_t pidx = first_index + p;
        unsigned long long best = ~0ull;
        for (int sb = s0; sb < s1; sb += 32) {
            const int s = sb + lane;
            const bool active = s < s1;
            float x0[D0];
            if (PHILOX) {
                if (grp == 4) fb_draw_spec<D0, 1, 4>(lo, range, s, pidx, seed, x0);
                else if (grp == 2) fb_draw_spec<D0, 1, 2>(lo, range, s, pidx, seed, x0);
                else fb_draw_spec<D0, 1, 1>(lo, range, s, pidx, seed, x0);
            } else {
                const int32_t *row = samples + ((int64_t)p * nsamp + (active ? s : s0)) * D0;
#pragma unroll
                for (int i = 0; i < D0; ++i) x0[i] = (float)__ldg(row + i);
            }
            bool havep = false, haven = false;
            int vp = 0, vn = 0;
            FbReluVisit<float> vis;
            for (int v = pa_lo; v <= pa_hi; ++v) {
                float x[D0];
#pragma unroll
                for (int i = 0; i < D0; ++i) x[i] = (i == pa_col) ? (float)v : x0[i];
                const float y = fb_spec_forward<S, 0, float>(prm, x, vis);
                if (y > 0.0f && !havep) { havep = true; vp = v; }
                if (y < 0.0f && !haven) { haven = true; vn = v; }
            }
            const bool ce = active && havep && haven;
            if (ce && best == ~0ull)
                best = ((unsigned long long)(uint32_t)s << 32) | ((unsigned long long)(uint32_t)(vp - pa_lo) << 16) |
                       (unsigned long long)(uint32_t)(vn - pa_lo);
            if (__any_sync(0xffffffffu, ce)) break;          // later samples cannot beat this step's
        }
        // lowest sample index wins; keys are unique per sample so min over the packed key is enough
        const uint32_t hi = (uint32_t)(best >> 32);
        const uint32_t hmin = __reduce_min_sync(0xffffffffu, hi);
        if (hmin != 0xffffffffu && hi == hmin) atomicMin(&ce_key[p], best);
    }
}
