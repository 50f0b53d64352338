// fb_spec_profile.cuh -- compile-time-specialised activation-profile kernel (see fb_common.cuh, fb_spec_fwd.cuh).
// The FP64 pipe is the only unit this kernel leans on: per sample P DFMA + N DADD.
#pragma once
#include "fb_spec_fwd.cuh"

// ---------------------------------------------------------------------------------------------------
// Activation profile.  reference: candidate_dead_nodes, utils/prune.py:168-192 (activation != 0 counted per
// neuron, output layer included) over the samples of simluate_data (:205-222).
//
// One warp per (partition, sample chunk).  Each lane carries K samples per step (VecD<K>: 2 where the register
// budget allows, else 1), so a step covers 32*K samples and every weight fetched from the constant bank feeds K
// fused multiply-adds.  The network itself is fb_spec_forward (layer-pair fused, terms in ascending input order).
// ---------------------------------------------------------------------------------------------------
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

