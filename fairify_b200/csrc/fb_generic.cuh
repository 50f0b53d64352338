// fb_generic.cuh -- runtime-shape kernels (see fb_common.cuh).  Same arithmetic, same term order as
// the specialised kernels and the oracle; used for networks outside FB_SPEC_LIST.
#pragma once
#include "fb_common.cuh"
#include "fb_spec.cuh"   // FbPruneOut, FbPaCols, thread-count macros

// ---------------------------------------------------------------------------------------------------
// forward pass, one sample per thread, activations in two local arrays, outputs blocked by 4 so one
// activation load feeds four independent FMA chains.  Visit::neuron(idx, last, pre) -> activation.
// ---------------------------------------------------------------------------------------------------
template <typename T, class Visit>
__device__ __forceinline__ T fb_gen_forward(const GenNet &net, const T *__restrict__ W, const T *__restrict__ B,
                                            T *in, T *out, Visit &vis)
{
    for (int l = 0; l < net.L; ++l) {
        const int IN = net.dims[l], OUT = net.dims[l + 1];
        const T *w = W + net.woff[l];
        const T *b = B + net.boff[l];
        const bool last = (l == net.L - 1);
        const int full = OUT & ~3;
        int j0 = 0;
        for (; j0 < full; j0 += 4) {
            T a0 = (T)0, a1 = (T)0, a2 = (T)0, a3 = (T)0;
            const T *wr = w + j0;
            for (int i = 0; i < IN; ++i, wr += OUT) {
                const T xv = in[i];
                a0 = fb_fma(wr[0], xv, a0); a1 = fb_fma(wr[1], xv, a1);
                a2 = fb_fma(wr[2], xv, a2); a3 = fb_fma(wr[3], xv, a3);
            }
            out[j0 + 0] = vis.neuron(net.boff[l] + j0 + 0, last, fb_add(a0, b[j0 + 0]));
            out[j0 + 1] = vis.neuron(net.boff[l] + j0 + 1, last, fb_add(a1, b[j0 + 1]));
            out[j0 + 2] = vis.neuron(net.boff[l] + j0 + 2, last, fb_add(a2, b[j0 + 2]));
            out[j0 + 3] = vis.neuron(net.boff[l] + j0 + 3, last, fb_add(a3, b[j0 + 3]));
        }
        for (; j0 < OUT; ++j0) {
            T a0 = (T)0;
            const T *wr = w + j0;
            for (int i = 0; i < IN; ++i, wr += OUT) a0 = fb_fma(wr[0], in[i], a0);
            out[j0] = vis.neuron(net.boff[l] + j0, last, fb_add(a0, b[j0]));
        }
        T *t = in; in = out; out = t;
    }
    return in[0];
}

template <typename T>
struct FbGenReluVisit {
    __device__ __forceinline__ T neuron(int, bool last, T pre) { return last ? pre : (fb_pos(pre) ? pre : (T)0); }
};

template <typename T>
struct FbGenMaskedVisit {
    const uint32_t *dead;     // [NW] or nullptr
    __device__ __forceinline__ T neuron(int idx, bool last, T pre)
    {
        if (last) return pre;
        bool keep = fb_pos(pre);
        if (dead) keep = keep && !((dead[idx >> 5] >> (idx & 31)) & 1u);
        return keep ? pre : (T)0;
    }
};

// warp-cooperative counting: every lane of the warp reaches neuron() together (uniform control flow)
struct FbGenCountVisit {
    uint32_t *cnt;     // this warp's [N] counters in shared memory
    bool active;
    int lane;
    __device__ __forceinline__ double neuron(int idx, bool last, double pre)
    {
        const bool nz = last ? fb_nonzero(pre) : fb_pos(pre);
        const uint32_t bal = __ballot_sync(0xffffffffu, nz && active);
        if (lane == (idx & 31)) cnt[idx] += __popc(bal);
        return last ? pre : (fb_pos(pre) ? pre : 0.0);
    }
};

// ---------------------------------------------------------------------------------------------------
// Activation profile for wide runtime-shape networks, organised like a tiled matrix product.
//
// A CTA takes TM samples of one partition through the whole network.  A layer is  C[m][n] = sum_k A[m][k] W[k][n]
// with the samples as rows: the activations of the TM samples sit in shared memory ([k][m], two buffers that swap
// roles from layer to layer), a warp owns a 32-sample x 32-neuron block of C, and a thread an (4 samples) x (8
// neurons) register tile whose 32 accumulators each run  acc = fma(w, x, acc)  over k ascending -- per neuron exactly
// the chain of the oracle and of the other kernels, so the counts are bit-identical.  Per k a thread loads 4
// activations (shared memory, the 4 lanes that share the samples read the same words) and 8 weights (shared memory
// too: the layer's rows, padded to 32 columns, stream through three cp.async stages of 16 rows) for 32 multiply-adds,
// against one activation load per multiply-add of the sample-per-lane kernel below.
// counts[p][N] is atomically added to (caller zeroes first).
// ---------------------------------------------------------------------------------------------------
struct FbTileNet {
    int L, N;
    int dims[FB_MAX_LAYERS + 1];     // true widths
    int dimsp[FB_MAX_LAYERS + 1];    // widths rounded up to 32 (row pitch of the padded weights)
    int woffp[FB_MAX_LAYERS + 1];    // layer offsets into Wp
    int boff[FB_MAX_LAYERS + 1];
    int maxw;                        // widest layer (inputs included): pitch of the activation buffers is TM
    const double *Wp;                // [in][dimsp[l+1]] per layer, zero padded
    const double *B;
    const float *Wp32;               // the same in float32 (partition screen)
    const float *Wiv32;              // rows (max(w_k, 0), min(w_k, 0)) interleaved: [2 in][dimsp[l+1]] per layer at 2 * woffp[l]
    const float *B32;
};

#define FB_TILE_THREADS 256
#define FB_TILE_KC 16            // weight rows per shared-memory stage

__device__ __forceinline__ void fb_cp_async16(void *smem_dst, const void *gsrc)
{
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void fb_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void fb_cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

#define FB_TILE_STAGES 3         // weight stages in flight: with three, one barrier per stage suffices
// dynamic shared memory: two activation buffers [maxw][TM], the weight stages [KC][maxwp], per-neuron counters [N]
__host__ __device__ inline size_t fb_tile_smem_bytes(int maxw, int maxwp, int N, int TM)
{
    return sizeof(double) * (2 * (size_t)maxw * TM + (size_t)FB_TILE_STAGES * FB_TILE_KC * maxwp) + sizeof(uint32_t) * N;
}
// ... with all `wdoubles` padded weights resident instead of the stages
__host__ __device__ inline size_t fb_tile_smem_bytes_resident(int maxw, int wdoubles, int N, int TM)
{
    return sizeof(double) * (2 * (size_t)maxw * TM + (size_t)wdoubles) + sizeof(uint32_t) * N;
}

// RESIDENT: the whole padded weight array fits in shared memory beside the activations and is loaded once per CTA;
// layers then run without stages and without the per-stage barriers.
template <int TM, bool PHILOX, bool RESIDENT>
__global__ void __launch_bounds__(FB_TILE_THREADS)
fb_profile_tile_kernel(const FbTileNet net, int64_t n, const int32_t *__restrict__ boxes,
                       const int32_t *__restrict__ samples, uint64_t seed, int64_t first_index, int nsamp,
                       uint32_t *__restrict__ counts)
{
    extern __shared__ __align__(16) unsigned char fb_smem[];
    constexpr int KC = FB_TILE_KC;
    double *buf0 = reinterpret_cast<double *>(fb_smem);               // activations: two [maxw][TM] buffers
    const int bufsz = net.maxw * TM;
    const int maxwp = (net.maxw + 31) & ~31;
    double *wst = buf0 + 2 * (size_t)bufsz;                            // weights: FB_TILE_STAGES stages of [KC][maxwp], or all of them
    const size_t wdoubles = RESIDENT ? (size_t)net.woffp[net.L] : (size_t)FB_TILE_STAGES * KC * maxwp;
    uint32_t *scnt = reinterpret_cast<uint32_t *>(wst + wdoubles);
    if (RESIDENT) {
        const double2 *src = reinterpret_cast<const double2 *>(net.Wp);
        double2 *dst = reinterpret_cast<double2 *>(wst);
        for (int i = threadIdx.x; i < net.woffp[net.L] / 2; i += FB_TILE_THREADS) dst[i] = __ldg(src + i);
    }
    constexpr int MB = TM / 32;                           // 32-sample blocks of the tile
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int lane_m = lane & 7, lane_n = lane >> 3;      // a warp's block: 8 lanes x 4 samples by 4 lanes x 8 neurons
    const int D0 = net.dims[0], N = net.N;
    const int tiles = (nsamp + TM - 1) / TM;
    const int64_t units = n * tiles;
    for (int k = tid; k < N; k += FB_TILE_THREADS) scnt[k] = 0;
    for (int64_t u = blockIdx.x; u < units; u += gridDim.x) {
        const int64_t p = u / tiles;
        const int s0 = (int)(u - p * tiles) * TM;
        const int live = min(TM, nsamp - s0);             // samples of this tile that exist
        // -- this tile's samples, feature-major: buf0[i * TM + m]
        if (tid < TM) {
            double *col = buf0 + tid;
            if (tid < live) {
                if (PHILOX) {
                    const int32_t *box = boxes + p * (2 * D0);
                    fb_draw_row<double>(box, D0, fb_draw_group(fb_box_widest(box, D0)), (uint32_t)(s0 + tid),
                                        first_index + p, seed, col, TM);
                } else {
                    const int32_t *row = samples + ((int64_t)p * nsamp + s0 + tid) * D0;
                    for (int i = 0; i < D0; ++i) col[(int64_t)i * TM] = (double)__ldg(row + i);
                }
            } else {
                for (int i = 0; i < D0; ++i) col[(int64_t)i * TM] = 0.0;
            }
        }
        __syncthreads();
        for (int l = 0; l < net.L; ++l) {
            const int IN = net.dims[l], OUT = net.dims[l + 1], OUTP = net.dimsp[l + 1];
            const bool last = (l == net.L - 1);
            const double *Ain = buf0 + (l & 1) * bufsz;
            double *Aout = buf0 + ((l + 1) & 1) * bufsz;
            const int nblocks = MB * (OUTP / 32);
            const int nchunks = RESIDENT ? 1 : (IN + KC - 1) / KC;
            const double *Wl = net.Wp + net.woffp[l];
            if (OUT <= 8) {
                // narrow layer (the logit, the tail of a funnel): a 32-column block would be mostly padding, so every
                // thread takes whole dot products instead -- (sample, neuron) pairs, a warp's 32 pairs share the neuron
                for (int t = tid; t < TM * OUT; t += FB_TILE_THREADS) {
                    const int mm = t % TM, nn = t / TM;
                    const double *ap = Ain + mm;
                    const double *wp = (RESIDENT ? wst + net.woffp[l] : Wl) + nn;
                    double acc = 0.0;
#pragma unroll 4
                    for (int k = 0; k < IN; ++k) acc = __fma_rn(wp[(size_t)k * OUTP], ap[(size_t)k * TM], acc);
                    const double pre = __dadd_rn(acc, __ldg(net.B + net.boff[l] + nn));
                    const bool pos = fb_pos(pre);
                    const bool nz = (last ? fb_nonzero(pre) : pos) && (mm < live);
                    const uint32_t bal = __ballot_sync(0xffffffffu, nz);
                    if (lane == 0 && bal) atomicAdd(&scnt[net.boff[l] + nn], (uint32_t)__popc(bal));
                    if (!last) Aout[(size_t)nn * TM + mm] = pos ? pre : 0.0;
                }
                __syncthreads();
                continue;
            }
            // the layer's weights stream through shared memory KC rows at a time (cp.async, two stages), all columns
            auto stage = [&](int c) {
                const int rows = min(KC, IN - c * KC);
                const double2 *src = reinterpret_cast<const double2 *>(Wl + (size_t)c * KC * OUTP);
                double2 *dst = reinterpret_cast<double2 *>(wst + (size_t)(c % FB_TILE_STAGES) * KC * maxwp);
                for (int i = tid; i < rows * (OUTP / 2); i += FB_TILE_THREADS) fb_cp_async16(dst + i, src + i);
                fb_cp_async_commit();
            };
            for (int blk0 = 0; blk0 < nblocks; blk0 += FB_TILE_THREADS / 32) {       // rounds: one block per warp
                const int blk = blk0 + warp;
                const bool active = blk < nblocks;
                const int mb = blk % MB, nb = blk / MB;
                // this thread's 4 samples: two adjacent pairs 16 apart, m0 + {0, 1, 16, 17} -- a quarter-warp then reads
                // 128 contiguous bytes per 128-bit load (no bank conflicts); its 8 neurons: n0 .. n0 + 7
                const int m0 = mb * 32 + lane_m * 2, n0 = nb * 32 + lane_n * 8;
                double acc[4][8];
#pragma unroll
                for (int r = 0; r < 4; ++r)
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[r][j] = 0.0;
                // stage c is consumed in iteration c while c+1 and c+2 are in flight; the barrier at the top of an
                // iteration also says that everybody is done with stage c-1, whose slot stage c+2 is about to overwrite
                if (!RESIDENT) {
                    stage(0);
                    if (nchunks > 1) stage(1);
                }
                for (int c = 0; c < nchunks; ++c) {
                    if (!RESIDENT) {
                        if (c + 1 < nchunks) fb_cp_async_wait<1>();
                        else fb_cp_async_wait<0>();
                        __syncthreads();
                        if (c + 2 < nchunks) stage(c + 2);
                    }
                    if (active) {
                        const int rows = RESIDENT ? IN : min(KC, IN - c * KC);
                        const double *ap = Ain + (size_t)c * KC * TM + m0;
                        const double *wp = RESIDENT ? wst + net.woffp[l] + n0 : wst + (size_t)(c % FB_TILE_STAGES) * KC * maxwp + n0;
                        // operands of row k+1 are fetched while row k's multiply-adds issue (two register sets, A / B)
                        double2 xa[2], wa[4], xb[2], wb[4];
#define FB_TILE_LOAD(X, W, AP, WP)                                                                         \
    {                                                                                                       \
        X[0] = *reinterpret_cast<const double2 *>(AP); X[1] = *reinterpret_cast<const double2 *>((AP) + 16); \
        W[0] = reinterpret_cast<const double2 *>(WP)[0]; W[1] = reinterpret_cast<const double2 *>(WP)[1];   \
        W[2] = reinterpret_cast<const double2 *>(WP)[2]; W[3] = reinterpret_cast<const double2 *>(WP)[3];   \
    }
#define FB_TILE_FMA(X, W)                                                                                  \
    {                                                                                                       \
        const double a_[4] = {X[0].x, X[0].y, X[1].x, X[1].y};                                              \
        const double w_[8] = {W[0].x, W[0].y, W[1].x, W[1].y, W[2].x, W[2].y, W[3].x, W[3].y};              \
        _Pragma("unroll") for (int r = 0; r < 4; ++r)                                                       \
            _Pragma("unroll") for (int j = 0; j < 8; ++j) acc[r][j] = __fma_rn(w_[j], a_[r], acc[r][j]);    \
    }
                        FB_TILE_LOAD(xa, wa, ap, wp);
                        for (int k = 0; k < rows; k += 2) {
                            const bool odd = k + 1 < rows, more = k + 2 < rows;
                            if (odd) FB_TILE_LOAD(xb, wb, ap + TM, wp + OUTP);
                            FB_TILE_FMA(xa, wa);
                            ap += 2 * TM; wp += 2 * OUTP;
                            if (odd) {
                                if (more) FB_TILE_LOAD(xa, wa, ap, wp);
                                FB_TILE_FMA(xb, wb);
                            }
                        }
#undef FB_TILE_LOAD
#undef FB_TILE_FMA
                    }
                }
                if (!RESIDENT) __syncthreads();           // the next round (or layer) restarts the stages
                if (!active) continue;
                // bias, ReLU, non-zero counts; hidden activations go to the other buffer
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int nn = n0 + j;
                    const bool valid = nn < OUT;
                    const double bj = valid ? __ldg(net.B + net.boff[l] + nn) : 0.0;
                    uint32_t cnt = 0;
                    double h[4];
#pragma unroll
                    for (int r = 0; r < 4; ++r) {
                        const double pre = __dadd_rn(acc[r][j], bj);
                        const bool pos = fb_pos(pre);
                        const bool nz = (last ? fb_nonzero(pre) : pos) && (m0 + (r >> 1) * 16 + (r & 1) < live) && valid;
                        const uint32_t bal = __ballot_sync(0xffffffffu, nz);
                        cnt += __popc(bal & (0xffu << (lane_n * 8)));
                        h[r] = pos ? pre : 0.0;
                    }
                    if (valid) {
                        if (lane_m == 0 && cnt) atomicAdd(&scnt[net.boff[l] + nn], cnt);
                        if (!last) {
                            double *op = Aout + (size_t)nn * TM + m0;
                            *reinterpret_cast<double2 *>(op) = make_double2(h[0], h[1]);
                            *reinterpret_cast<double2 *>(op + 16) = make_double2(h[2], h[3]);
                        }
                    }
                }
            }
            __syncthreads();
        }
        for (int k = tid; k < N; k += FB_TILE_THREADS) {
            const uint32_t c = scnt[k];
            if (c) atomicAdd(&counts[p * N + k], c);
            scnt[k] = 0;
        }
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------------------
// Per-partition sampled counterexample screen for wide runtime-shape networks, float32, organised like the tiled
// profile kernel above: the rows of the product are (sample, protected value) pairs.  One CTA per partition walks
// the samples in tiles of TM rows and stops at the first tile that holds a counterexample, so the answer -- the
// lowest sample index whose sweep of the protected column produces both classes, with the first positive and the
// first negative value -- is the one the sample-per-lane kernel gives.  A sweep wider than a tile is cut into
// segments of TM values of one sample.
// ---------------------------------------------------------------------------------------------------
__host__ __device__ inline size_t fb_screen_tile_smem_bytes(int maxw, int maxwp, int TM)
{
    // activations (two buffers), weight stages, the best key + state words, the relaxed screen's pending-sample list
    return sizeof(float) * (2 * (size_t)maxw * TM + (size_t)FB_TILE_STAGES * FB_TILE_KC * maxwp) + 64 + sizeof(int) * (2 * (size_t)TM);
}

// all layers for the TM rows in buf0 (feature-major, `bufsz` floats per buffer); returns the buffer whose first feature
// row holds the TM logits.  NJ = neurons per thread of the 4 x NJ register tile (8, or 4 to spread a small tile over all
// warps).
//
// IV: the interval form.  Rows come in pairs (2s, 2s+1) = (lower, upper) end of sample s, and a layer with IN inputs
// is a plain product over 2 IN features: the lower row holds (lo_0, hi_0, lo_1, hi_1, ...), the upper row
// (hi_0, lo_0, hi_1, lo_1, ...), and the weight rows are (max(w_0, 0), min(w_0, 0), max(w_1, 0), ...) (Wiv32).  Of the two
// terms of input k one is the product the ordinary pass would add (lower end: w_k lo_k if w_k >= 0, else w_k hi_k) and
// the other is 0 * x, which leaves the sum as it is: each end is the same chain of round-to-nearest multiply-adds,
// in the same order, as the ordinary pass -- so [lower, upper] encloses every logit that pass COMPUTES for inputs
// between the ends (rounding is monotone).  The epilogue writes each neuron's two ends back in both orders.
template <int TM, int NJ, bool IV>
__device__ __forceinline__ const float *fb_tile_forward_f32(const FbTileNet &net, float *buf0, int bufsz, float *wst, int maxwp)
{
    constexpr int KC = FB_TILE_KC, MB = TM / 32, NB = 4 * NJ;           // a warp's block: 32 rows x NB neurons
    static_assert(NJ == 4 || NJ == 8, "4 x 4 or 4 x 8 register tiles");
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int lane_m = lane & 7, lane_n = lane >> 3;
    for (int l = 0; l < net.L; ++l) {
        const int IN = (IV ? 2 : 1) * net.dims[l], OUT = net.dims[l + 1], OUTP = net.dimsp[l + 1];
        const bool last = (l == net.L - 1);
        const float *Ain = buf0 + (l & 1) * bufsz;
        float *Aout = buf0 + ((l + 1) & 1) * bufsz;
        const float *Wl = IV ? net.Wiv32 + 2 * (size_t)net.woffp[l] : net.Wp32 + net.woffp[l];
        if (OUT <= 8) {
            for (int t = tid; t < TM * OUT; t += FB_TILE_THREADS) {
                const int mm = t % TM, nn = t / TM;
                const float *ap = Ain + mm;
                const float *wp = Wl + nn;
                float acc = 0.0f;
#pragma unroll 4
                for (int k = 0; k < IN; ++k) acc = __fmaf_rn(__ldg(wp + (size_t)k * OUTP), ap[(size_t)k * TM], acc);
                const float pre = __fadd_rn(acc, __ldg(net.B32 + net.boff[l] + nn));
                const float h = last ? pre : (fb_pos(pre) ? pre : 0.0f);
                if (IV && !last) {
                    Aout[(size_t)(2 * nn) * TM + mm] = h;                 // own end first in the own row ...
                    Aout[(size_t)(2 * nn + 1) * TM + (mm ^ 1)] = h;       // ... and second in the partner row
                } else {
                    Aout[(size_t)nn * TM + mm] = h;
                }
            }
            __syncthreads();
            continue;
        }
        const int nblocks = MB * (OUTP / NB);
        const int nchunks = (IN + KC - 1) / KC;
        auto stage = [&](int c) {
            const int rows = min(KC, IN - c * KC);
            const float4 *src = reinterpret_cast<const float4 *>(Wl + (size_t)c * KC * OUTP);
            float4 *dst = reinterpret_cast<float4 *>(wst + (size_t)(c % FB_TILE_STAGES) * KC * maxwp);
            for (int i = tid; i < rows * (OUTP / 4); i += FB_TILE_THREADS) fb_cp_async16(dst + i, src + i);
            fb_cp_async_commit();
        };
        for (int blk0 = 0; blk0 < nblocks; blk0 += FB_TILE_THREADS / 32) {
            const int blk = blk0 + warp;
            const bool active = blk < nblocks;
            const int mb = blk % MB, nb = blk / MB;
            const int m0 = mb * 32 + lane_m * 4, n0 = nb * NB + lane_n * NJ;      // 4 adjacent rows, NJ adjacent neurons
            float acc[4][NJ];
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int j = 0; j < NJ; ++j) acc[r][j] = 0.0f;
            stage(0);
            if (nchunks > 1) stage(1);
            for (int c = 0; c < nchunks; ++c) {
                if (c + 1 < nchunks) fb_cp_async_wait<1>();
                else fb_cp_async_wait<0>();
                __syncthreads();
                if (c + 2 < nchunks) stage(c + 2);
                if (active) {
                    const int rows = min(KC, IN - c * KC);
                    const float *ap = Ain + (size_t)c * KC * TM + m0;
                    const float *wp = wst + (size_t)(c % FB_TILE_STAGES) * KC * maxwp + n0;
#pragma unroll 4
                    for (int k = 0; k < rows; ++k, ap += TM, wp += OUTP) {
                        const float4 a4 = *reinterpret_cast<const float4 *>(ap);
                        const float a[4] = {a4.x, a4.y, a4.z, a4.w};
                        float w[NJ];
#pragma unroll
                        for (int q = 0; q < NJ / 4; ++q) {
                            const float4 w4 = reinterpret_cast<const float4 *>(wp)[q];
                            w[4 * q] = w4.x; w[4 * q + 1] = w4.y; w[4 * q + 2] = w4.z; w[4 * q + 3] = w4.w;
                        }
#pragma unroll
                        for (int r = 0; r < 4; ++r)
#pragma unroll
                            for (int j = 0; j < NJ; ++j) acc[r][j] = __fmaf_rn(w[j], a[r], acc[r][j]);
                    }
                }
            }
            __syncthreads();
            if (!active) continue;
#pragma unroll
            for (int j = 0; j < NJ; ++j) {
                const int nn = n0 + j;
                if (nn < OUT) {
                    const float bj = __ldg(net.B32 + net.boff[l] + nn);
                    float h[4];
#pragma unroll
                    for (int r = 0; r < 4; ++r) {
                        const float pre = __fadd_rn(acc[r][j], bj);
                        h[r] = last ? pre : (fb_pos(pre) ? pre : 0.0f);
                    }
                    if (IV && !last) {
                        *reinterpret_cast<float4 *>(Aout + (size_t)(2 * nn) * TM + m0) = make_float4(h[0], h[1], h[2], h[3]);
                        *reinterpret_cast<float4 *>(Aout + (size_t)(2 * nn + 1) * TM + m0) = make_float4(h[1], h[0], h[3], h[2]);
                    } else {
                        *reinterpret_cast<float4 *>(Aout + (size_t)nn * TM + m0) = make_float4(h[0], h[1], h[2], h[3]);
                    }
                }
            }
        }
        __syncthreads();
    }
    return buf0 + (net.L & 1) * bufsz;
}

template <int TM, bool PHILOX>
__global__ void __launch_bounds__(FB_TILE_THREADS)
fb_screen_tile_kernel(const FbTileNet net, int64_t n, const int32_t *__restrict__ boxes,
                      const int32_t *__restrict__ samples, uint64_t seed, int64_t first_index, int nsamp, int pa_col,
                      int ra_col, int ra_thr, unsigned long long *ce_key)
{
    extern __shared__ __align__(16) unsigned char fb_smem[];
    float *buf0 = reinterpret_cast<float *>(fb_smem);
    const int bufsz = net.maxw * TM;
    const int maxwp = (net.maxw + 31) & ~31;
    float *wst = buf0 + 2 * (size_t)bufsz;
    unsigned long long *s_best = reinterpret_cast<unsigned long long *>(wst + (size_t)FB_TILE_STAGES * FB_TILE_KC * maxwp);
    int *s_state = reinterpret_cast<int *>(s_best + 1);                 // "this sample's sweep is decided" of a segmented sweep
    const int tid = threadIdx.x;
    const int D0 = net.dims[0];
    const bool relaxed = ra_col >= 0;
    const int nd = relaxed ? 2 * ra_thr + 1 : 1;                        // RA offsets per protected value
    for (int64_t p = blockIdx.x; p < n; p += gridDim.x) {
        if (ce_key[p] == FB_SCREEN_NOSEARCH) continue;                   // (block-uniform: only this CTA touches ce_key[p])
        const int32_t *box = boxes + p * (2 * D0);
        const int pa_lo = __ldg(box + 2 * pa_col), pa_hi = __ldg(box + 2 * pa_col + 1);
        const uint32_t Wu = (uint32_t)(pa_hi - pa_lo + 1);
        if (Wu == 0 || Wu > (relaxed ? FB_SCREEN_MAX_W_RELAXED : FB_SCREEN_MAX_W_STRICT)) {   // does not fit the key's fields
            if (tid == 0) ce_key[p] = FB_SCREEN_SKIPPED;
            continue;
        }
        const int W = (int)Wu;
        const int R = W * nd;                                            // rows of one sample's sweep
        const int grp = PHILOX ? fb_draw_group(fb_box_widest(box, D0)) : 1;
        const int64_t pidx = first_index + p;
        if (tid == 0) *s_best = ~0ull;
        __syncthreads();
        // row t of a tile: sample s_row, row q of its sweep (or a zero row past the end)
        auto fill = [&](int s_row, int q, bool exists) {
            float *col = buf0 + tid;
            if (exists) {
                if (PHILOX) fb_draw_row<float>(box, D0, grp, (uint32_t)s_row, pidx, seed, col, TM);
                else {
                    const int32_t *row = samples + ((int64_t)p * nsamp + s_row) * D0;
                    for (int i = 0; i < D0; ++i) col[(size_t)i * TM] = (float)__ldg(row + i);
                }
                const int k = q / W, off = q - k * W;
                col[(size_t)pa_col * TM] = (float)(pa_lo + off);
                if (relaxed) {
                    int d = k == 0 ? 0 : -ra_thr + (k - 1);
                    if (k != 0 && d >= 0) ++d;
                    col[(size_t)ra_col * TM] = (float)((int)col[(size_t)ra_col * TM] + d);
                }
            } else {
                for (int i = 0; i < D0; ++i) col[(size_t)i * TM] = 0.0f;
            }
        };
        if (R <= TM && relaxed) {
            // Relaxed sweeps are W * (2 thr + 1) rows per sample, and a box without a pair has every sample swept to
            // the end.  So each sample first goes through the network once as an INTERVAL (two rows: protected column
            // over the box's range, relaxed column over a - thr .. a + thr, the other features at the sample's values;
            // fb_tile_forward_f32<.., true>).  A one-signed enclosure of the logit means the sweep would find one class
            // only.  The samples left over wait in `pend`, ascending, until they fill a tile of sweeps.
            constexpr int TA = TM / 2, SA = TA / 2;                       // interval tile: TA rows = SA samples
            const int spt = TM / R;                                       // whole sweeps per tile
            int *pend = s_state + 4;                                      // up to spt - 1 + SA entries
            int npend = 0, sb = 0;
            while (sb < nsamp || npend > 0) {
                if (sb < nsamp && npend < spt) {
                    const int ns = min(SA, nsamp - sb);
                    if (tid < TA) {
                        float *col = buf0 + tid;
                        const int s_row = sb + (tid >> 1);
                        const bool upper = tid & 1;
                        if ((tid >> 1) < ns) {
                            if (PHILOX) fb_draw_row<float>(box, D0, grp, (uint32_t)s_row, pidx, seed, col, 2 * TA);
                            else {
                                const int32_t *row = samples + ((int64_t)p * nsamp + s_row) * D0;
                                for (int i = 0; i < D0; ++i) col[(size_t)(2 * i) * TA] = (float)__ldg(row + i);
                            }
                            for (int i = 0; i < D0; ++i) {
                                float lo_i = col[(size_t)(2 * i) * TA], hi_i = lo_i;
                                if (i == pa_col) { lo_i = (float)pa_lo; hi_i = (float)pa_hi; }
                                if (i == ra_col) { const int a = (int)lo_i; lo_i = (float)(a - ra_thr); hi_i = (float)(a + ra_thr); }
                                col[(size_t)(2 * i) * TA] = upper ? hi_i : lo_i;
                                col[(size_t)(2 * i + 1) * TA] = upper ? lo_i : hi_i;
                            }
                        } else {
                            for (int i = 0; i < 2 * D0; ++i) col[(size_t)i * TA] = 0.0f;
                        }
                    }
                    __syncthreads();
                    const float *y = fb_tile_forward_f32<TA, 4, true>(net, buf0, bufsz, wst, maxwp);
                    if (tid == 0) {
                        int np = npend;
                        for (int t = 0; t < ns; ++t)
                            if (!(y[2 * t] > 0.0f || y[2 * t + 1] < 0.0f)) pend[np++] = sb + t;
                        s_state[0] = np;
                    }
                    __syncthreads();
                    npend = s_state[0];
                    sb += SA;
                    continue;
                }
                const int take = min(npend, spt);
                if (tid < TM) {
                    const bool exists = tid < take * R;
                    fill(exists ? pend[tid / R] : 0, tid % R, exists);
                }
                __syncthreads();
                const float *y = fb_tile_forward_f32<TM, 8, false>(net, buf0, bufsz, wst, maxwp);
                if (tid < take) {
                    FbScreenScan sc;
                    sc.begin();
                    for (int q = 0; q < R && !sc.done; ++q) sc.row(q, y[tid * R + q], W, relaxed, ra_thr);
                    if (sc.low != 0xffffffffu)
                        atomicMin(s_best, ((unsigned long long)(uint32_t)pend[tid] << 32) | (unsigned long long)sc.low);
                }
                __syncthreads();
                if (*s_best != ~0ull) break;
                const int rest = npend - take;
                const int v = tid < rest ? pend[take + tid] : 0;
                __syncthreads();
                if (tid < rest) pend[tid] = v;
                __syncthreads();
                npend = rest;
            }
        } else if (R <= TM) {
            const int spt = TM / R;                                       // whole samples per tile
            for (int sbase = 0; sbase < nsamp; sbase += spt) {
                const int ns = min(spt, nsamp - sbase);
                if (tid < TM) fill(sbase + tid / R, tid % R, tid < ns * R);
                __syncthreads();
                const float *y = fb_tile_forward_f32<TM, 8, false>(net, buf0, bufsz, wst, maxwp);
                if (tid < ns) {
                    FbScreenScan sc;
                    sc.begin();
                    for (int q = 0; q < R && !sc.done; ++q) sc.row(q, y[tid * R + q], W, relaxed, ra_thr);
                    if (sc.low != 0xffffffffu)
                        atomicMin(s_best, ((unsigned long long)(uint32_t)(sbase + tid) << 32) | (unsigned long long)sc.low);
                }
                __syncthreads();
                if (*s_best != ~0ull) break;
            }
        } else {
            FbScreenScan sc;                                              // thread 0's, carried across the segments
            for (int s = 0; s < nsamp; ++s) {
                sc.begin();
                if (tid == 0) s_state[0] = 0;
                for (int q0 = 0; q0 < R; q0 += TM) {
                    const int nq = min(TM, R - q0);
                    if (tid < TM) fill(s, q0 + tid, tid < nq);
                    __syncthreads();
                    const float *y = fb_tile_forward_f32<TM, 8, false>(net, buf0, bufsz, wst, maxwp);
                    if (tid == 0) {
                        for (int t = 0; t < nq && !sc.done; ++t) sc.row(q0 + t, y[t], W, relaxed, ra_thr);
                        if (sc.done) {
                            s_state[0] = 1;
                            if (sc.low != 0xffffffffu) *s_best = ((unsigned long long)(uint32_t)s << 32) | (unsigned long long)sc.low;
                        }
                    }
                    __syncthreads();
                    if (s_state[0]) break;
                }
                __syncthreads();
                if (*s_best != ~0ull) break;
            }
        }
        if (tid == 0 && *s_best != ~0ull) ce_key[p] = *s_best;
        __syncthreads();
    }
}

// weights (float64) staged once per CTA into dynamic shared memory: [P] weights, [N] biases, then
// per-warp uint32 counters [warps][N].
template <bool PHILOX>
__global__ void __launch_bounds__(FB_PROFILE_THREADS)
fb_profile_gen_kernel(const GenNet net, int64_t n, const int32_t *__restrict__ boxes,
                      const int32_t *__restrict__ samples, uint64_t seed, int64_t first_index, int nsamp, int chunks,
                      int chunk_len, uint32_t *__restrict__ counts)
{
    extern __shared__ __align__(16) unsigned char fb_smem[];
    double *sW = reinterpret_cast<double *>(fb_smem);
    double *sB = sW + net.P;
    uint32_t *sC = reinterpret_cast<uint32_t *>(sB + net.N);
    for (int k = threadIdx.x; k < net.P; k += blockDim.x) sW[k] = net.W64[k];
    for (int k = threadIdx.x; k < net.N; k += blockDim.x) sB[k] = net.B64[k];
    __syncthreads();

    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int D0 = net.dims[0], N = net.N;
    uint32_t *cnt = sC + wib * N;
    const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
    const int64_t gwarp = (int64_t)blockIdx.x * (blockDim.x >> 5) + wib;
    const int64_t units = n * chunks;
    double a[FB_MAXW], y[FB_MAXW];
    for (int64_t u = gwarp; u < units; u += nwarps) {
        const int64_t p = u / chunks;
        const int c = (int)(u - p * chunks);
        const int s0 = c * chunk_len;
        const int s1 = min(nsamp, s0 + chunk_len);
        for (int k = lane; k < N; k += 32) cnt[k] = 0;
        __syncwarp();
        const int32_t *box = boxes + p * (2 * D0);
        const int64_t pidx = first_index + p;
        const int grp = PHILOX ? fb_draw_group(fb_box_widest(box, D0)) : 1;
        FbGenCountVisit vis{cnt, false, lane};
        for (int sb = s0; sb < s1; sb += 32) {
            const int s = sb + lane;
            vis.active = s < s1;
            if (PHILOX) {
                fb_draw_row<double>(box, D0, grp, (uint32_t)s, pidx, seed, a);
            } else {
                const int32_t *row = samples + ((int64_t)p * nsamp + (vis.active ? s : s0)) * D0;
                for (int i = 0; i < D0; ++i) a[i] = (double)__ldg(row + i);
            }
            fb_gen_forward<double>(net, sW, sB, a, y, vis);
        }
        __syncwarp();
        for (int k = lane; k < N; k += 32) {
            if (chunks == 1) counts[p * N + k] = cnt[k];
            else atomicAdd(&counts[p * N + k], cnt[k]);
        }
        __syncwarp();
    }
}

// ---------------------------------------------------------------------------------------------------
// interval recurrence, one warp per partition, one lane per neuron of the current layer; the layer's
// input bounds sit in shared memory (broadcast reads), weights stream from global memory coalesced
// across lanes (W[i][j], j = lane).
// ---------------------------------------------------------------------------------------------------
#define FB_IBP_GEN_WARPS 4

__global__ void __launch_bounds__(FB_IBP_GEN_WARPS * 32)
fb_ibp_gen_kernel(const GenNet net, int64_t n, const int32_t *__restrict__ boxes, const uint32_t *__restrict__ counts,
                  const double *__restrict__ tol, const FbPruneOut o)
{
    __shared__ double sb[FB_IBP_GEN_WARPS][8][FB_MAXW];      // lb, ub, lbd, ubd  x  {in, out}
    __shared__ uint32_t sm[FB_IBP_GEN_WARPS][7][FB_MAX_NW];  // cand, b_dead, sliver, b_cand, s_dead, s_cand, dead
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int N = net.N, NW = (N + 31) / 32, D0 = net.dims[0];
    const int64_t nwarps = (int64_t)gridDim.x * FB_IBP_GEN_WARPS;
    for (int64_t p = (int64_t)blockIdx.x * FB_IBP_GEN_WARPS + wib; p < n; p += nwarps) {
        double *lb = sb[wib][0], *ub = sb[wib][1], *lbd = sb[wib][2], *ubd = sb[wib][3];
        double *nlb = sb[wib][4], *nub = sb[wib][5], *nlbd = sb[wib][6], *nubd = sb[wib][7];
        uint32_t *cand_w = sm[wib][0], *bdead_w = sm[wib][1], *sliver_w = sm[wib][2], *bcand_w = sm[wib][3];
        uint32_t *sdead_w = sm[wib][4], *scand_w = sm[wib][5], *dead_w = sm[wib][6];
        for (int k = lane; k < NW; k += 32) {
            cand_w[k] = 0; bdead_w[k] = 0; sliver_w[k] = 0; bcand_w[k] = 0; sdead_w[k] = 0; scand_w[k] = 0; dead_w[k] = 0;
        }
        for (int i = lane; i < D0; i += 32) {
            const int lo = __ldg(boxes + p * (2 * D0) + 2 * i), hi = __ldg(boxes + p * (2 * D0) + 2 * i + 1);
            lb[i] = lbd[i] = (double)lo; ub[i] = ubd[i] = (double)hi;
        }
        __syncwarp();
        double out_lbd = 0.0, out_ubd = 0.0;
        for (int l = 0; l < net.L; ++l) {
            const int IN = net.dims[l], OUT = net.dims[l + 1], BO = net.boff[l];
            const double *w = net.W64 + net.woff[l];
            const bool last = (l == net.L - 1);
            const uint32_t first = 1u << (BO & 31);
            bool all1 = true, anyc = false;
            for (int jb = 0; jb < OUT; jb += 32) {
                const int j = jb + lane;
                const bool live = j < OUT;
                bool bd = true, bc = false;
                if (live) {
                    double mn = 0.0, mx = 0.0, mnd = 0.0, mxd = 0.0;
                    for (int i = 0; i < IN; ++i) {
                        const double wt = __ldg(w + (int64_t)i * OUT + j);
                        const bool neg = fb_signbit(wt);
                        const double a = neg ? ub[i] : lb[i], b = neg ? lb[i] : ub[i];
                        const double ad = neg ? ubd[i] : lbd[i], bdv = neg ? lbd[i] : ubd[i];
                        mn = __dadd_rn(mn, __dmul_rn(wt, a));
                        mx = __dadd_rn(mx, __dmul_rn(wt, b));
                        mnd = __dadd_rd(mnd, __dmul_rd(wt, ad));
                        mxd = __dadd_ru(mxd, __dmul_ru(wt, bdv));
                    }
                    const double bj = __ldg(net.B64 + BO + j);
                    const double wl = __dadd_rn(mn, bj), wu = __dadd_rn(mx, bj);
                    const double wld = __dadd_rd(mnd, bj), wud = __dadd_ru(mxd, bj);
                    const int idx = BO + j;
                    const uint32_t c32 = counts ? counts[p * N + idx] : 0u;
                    const bool c = c32 == 0, le0 = wu <= 0.0;
                    const uint32_t bit = 1u << (idx & 31);
                    if (c) atomicOr(&cand_w[idx >> 5], bit);
                    bd = last ? c : (c && le0);
                    bc = last ? c : (c && !le0);
                    if (bd) atomicOr(&bdead_w[idx >> 5], bit);
                    if (bc) atomicOr(&bcand_w[idx >> 5], bit);
                    if (l == 0 && !last && tol != nullptr && bc && wu <= tol[idx]) atomicOr(&sliver_w[idx >> 5], bit);
                    if (last) { out_lbd = wld; out_ubd = wud; }
                    if (o.counts16) o.counts16[p * N + idx] = (uint16_t)c32;
                    if (o.ws_lb) o.ws_lb[p * N + idx] = wl;
                    if (o.ws_ub) o.ws_ub[p * N + idx] = wu;
                    if (o.pl_lb) o.pl_lb[p * N + idx] = wl < 0.0 ? 0.0 : wl;
                    if (o.pl_ub) o.pl_ub[p * N + idx] = wu < 0.0 ? 0.0 : wu;
                    if (o.lb_rd) o.lb_rd[p * N + idx] = wld;
                    if (o.ub_ru) o.ub_ru[p * N + idx] = wud;
                    nlb[j] = wl < 0.0 ? 0.0 : wl; nub[j] = wu < 0.0 ? 0.0 : wu;
                    nlbd[j] = wld < 0.0 ? 0.0 : wld; nubd[j] = wud < 0.0 ? 0.0 : wud;
                }
                all1 = __all_sync(0xffffffffu, !live || bd) && all1;
                anyc = __any_sync(0xffffffffu, live && bc) || anyc;
            }
            __syncwarp();
            // singular verification in closed form (see FbIbpSink in fb_spec.cuh): hidden layers l >= 1, candidates the
            // bound did not settle; the inputs are this layer's own post-activation bounds nlb / nub
            bool sall1 = last;                                  // last layer: s_dead is the copied b_cand, then keep-one
            if (!last && l >= 1 && anyc) {
                const int MM = min(IN, OUT);
                sall1 = true;
                for (int jb = 0; jb < OUT; jb += 32) {
                    const int j = jb + lane, idx = BO + j;
                    const bool live = j < OUT;
                    bool sd = false;
                    if (live && ((bcand_w[idx >> 5] >> (idx & 31)) & 1u) && !((__ldg(net.unb + (idx >> 5)) >> (idx & 31)) & 1u)) {
                        double q = 0.0;
                        for (int k = 0; k < MM; ++k) {
                            const double wt = __ldg(w + (int64_t)k * OUT + j);
                            q = __dadd_rn(q, __dmul_rn(wt, fb_signbit(wt) ? nlb[k] : nub[k]));
                        }
                        q = __dadd_rn(q, __ldg(net.B64 + idx));
                        sd = q <= 0.0;
                        if (sd) atomicOr(&sdead_w[idx >> 5], 1u << (idx & 31));
                        if (tol != nullptr && fabs(q) <= tol[idx]) atomicOr(&sliver_w[idx >> 5], 1u << (idx & 31));
                    }
                    sall1 = __all_sync(0xffffffffu, !live || sd) && sall1;
                }
                __syncwarp();
            }
            // s_cand = b_cand minus what the query found; keep-one fixes on b_dead, s_dead and their union
            bool dall1 = true;
            for (int jb = 0; jb < OUT; jb += 32) {
                const int j = jb + lane, idx = BO + j;
                bool dd = true;
                if (j < OUT) {
                    const uint32_t bit = 1u << (idx & 31);
                    const bool bcb = bcand_w[idx >> 5] & bit, sdb = sdead_w[idx >> 5] & bit;
                    if (bcb && !sdb) atomicOr(&scand_w[idx >> 5], bit);
                    const bool clear_b = all1 && j == 0, clear_s = sall1 && j == 0;
                    dd = ((bdead_w[idx >> 5] & bit) && !clear_b) || (sdb && !clear_s);
                    if (dd) atomicOr(&dead_w[idx >> 5], bit);
                }
                dall1 = __all_sync(0xffffffffu, dd) && dall1;
            }
            __syncwarp();
            if (lane == 0) {
                if (all1) bdead_w[BO >> 5] &= ~first;
                if (sall1) sdead_w[BO >> 5] &= ~first;
                if (dall1) dead_w[BO >> 5] &= ~first;
            }
            double *t;
            t = lb; lb = nlb; nlb = t; t = ub; ub = nub; nub = t;
            t = lbd; lbd = nlbd; nlbd = t; t = ubd; ubd = nubd; nubd = t;
            __syncwarp();
        }
        // the output neuron is lane 0 of the last layer
        out_lbd = __shfl_sync(0xffffffffu, out_lbd, 0);
        out_ubd = __shfl_sync(0xffffffffu, out_ubd, 0);
        uint32_t anysl = 0;
        for (int k = lane; k < NW; k += 32) {
            if (o.cand) o.cand[p * NW + k] = cand_w[k];
            if (o.b_dead) o.b_dead[p * NW + k] = bdead_w[k];
            if (o.dead) o.dead[p * NW + k] = dead_w[k];
            if (o.sliver) o.sliver[p * NW + k] = sliver_w[k];
            if (o.s_dead) o.s_dead[p * NW + k] = sdead_w[k];
            if (o.s_cand) o.s_cand[p * NW + k] = scand_w[k];
            if (o.record) { o.record[p * (2 * NW + 2) + 1 + k] = dead_w[k]; o.record[p * (2 * NW + 2) + 1 + NW + k] = cand_w[k]; }
            anysl |= sliver_w[k];
        }
        anysl = __reduce_or_sync(0xffffffffu, anysl);
        if (lane == 0) {
            uint8_t v = 0;
            if (out_lbd > 0.0 || out_ubd < 0.0) v |= FB_VERDICT_CERTIFIED;
            if (anysl) v |= FB_VERDICT_SLIVER;
            if (o.verdict) o.verdict[p] = v;
            if (o.record) { o.record[p * (2 * NW + 2)] = v; o.record[p * (2 * NW + 2) + 1 + 2 * NW] = 0xffffffffu; }
        }
        __syncwarp();
    }
}

// ---------------------------------------------------------------------------------------------------
// logits / pair screen / partition screen, generic shapes: weights read through L1 (__ldg).
// ---------------------------------------------------------------------------------------------------
template <typename T> struct FbNetPtr;
template <> struct FbNetPtr<double> {
    static __device__ __forceinline__ const double *W(const GenNet &n) { return n.W64; }
    static __device__ __forceinline__ const double *B(const GenNet &n) { return n.B64; }
};
template <> struct FbNetPtr<float> {
    static __device__ __forceinline__ const float *W(const GenNet &n) { return n.W32; }
    static __device__ __forceinline__ const float *B(const GenNet &n) { return n.B32; }
};

template <typename T>
__global__ void __launch_bounds__(FB_FWD_THREADS)
fb_forward_gen_kernel(const GenNet net, int64_t m, const T *__restrict__ X, const uint32_t *__restrict__ dead_bits,
                      T *__restrict__ logits)
{
    const int D0 = net.dims[0];
    T a[FB_MAXW], y[FB_MAXW];
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < m; r += stride) {
        for (int i = 0; i < D0; ++i) a[i] = X[r * D0 + i];
        FbGenMaskedVisit<T> vis{dead_bits};
        logits[r] = fb_gen_forward<T>(net, FbNetPtr<T>::W(net), FbNetPtr<T>::B(net), a, y, vis);
    }
}

__global__ void __launch_bounds__(FB_FWD_THREADS)
fb_pair_screen_gen_kernel(const GenNet net, int64_t m, const float *__restrict__ X, const FbPaCols pa, int k_alts,
                          const float *__restrict__ alts, uint8_t *__restrict__ orig_pred, int32_t *__restrict__ first_flip)
{
    const int D0 = net.dims[0];
    float a[FB_MAXW], y[FB_MAXW];
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < m; r += stride) {
        const float *x0 = X + r * D0;
        for (int i = 0; i < D0; ++i) a[i] = x0[i];
        FbGenReluVisit<float> vis;
        const bool p0 = fb_pos(fb_gen_forward<float>(net, net.W32, net.B32, a, y, vis));
        int flip = -1;
        for (int al = 0; al < k_alts && flip < 0; ++al) {
            for (int i = 0; i < D0; ++i) a[i] = x0[i];
            bool same = true;
            for (int q = 0; q < pa.n; ++q) {
                const float v = __ldg(alts + al * pa.n + q);
                same = same && (v == x0[pa.col[q]]);
                a[pa.col[q]] = v;
            }
            if (same) continue;
            const bool p1 = fb_pos(fb_gen_forward<float>(net, net.W32, net.B32, a, y, vis));
            if (p1 != p0) flip = al;
        }
        orig_pred[r] = p0 ? 1 : 0;
        first_flip[r] = flip;
    }
}

template <bool PHILOX>
__global__ void __launch_bounds__(FB_PROFILE_THREADS)
fb_partition_screen_gen_kernel(const GenNet net, int64_t n, const int32_t *__restrict__ boxes,
                               const int32_t *__restrict__ samples, uint64_t seed, int64_t first_index, int nsamp,
                               int chunks, int chunk_len, int pa_col, int ra_col, int ra_thr,
                               unsigned long long *ce_key)
{
    const int D0 = net.dims[0];
    const int lane = threadIdx.x & 31;
    const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
    const int64_t gwarp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int64_t units = n * chunks;
    const bool relaxed = ra_col >= 0;
    float x0[FB_MAXW], a[FB_MAXW], y[FB_MAXW];
    for (int64_t u = gwarp; u < units; u += nwarps) {
        const int64_t p = u / chunks;
        if (ce_key[p] == FB_SCREEN_NOSEARCH) continue;
        const int c = (int)(u - p * chunks);
        const int s0 = c * chunk_len;
        const int s1 = min(nsamp, s0 + chunk_len);
        const int32_t *box = boxes + p * (2 * D0);
        const int pa_lo = __ldg(box + 2 * pa_col);
        const uint32_t W = (uint32_t)(__ldg(box + 2 * pa_col + 1) - pa_lo + 1);
        if (W == 0 || W > (relaxed ? FB_SCREEN_MAX_W_RELAXED : FB_SCREEN_MAX_W_STRICT)) {
            if (lane == 0 && c == 0) ce_key[p] = FB_SCREEN_SKIPPED;
            continue;
        }
        const int64_t pidx = first_index + p;
        const int grp = PHILOX ? fb_draw_group(fb_box_widest(box, D0)) : 1;
        unsigned long long best = ~0ull;
        for (int sb = s0; sb < s1; sb += 32) {
            const int s = sb + lane;
            const bool active = s < s1;
            if (PHILOX) {
                fb_draw_row<float>(box, D0, grp, (uint32_t)s, pidx, seed, x0);
            } else {
                const int32_t *row = samples + ((int64_t)p * nsamp + (active ? s : s0)) * D0;
                for (int i = 0; i < D0; ++i) x0[i] = (float)__ldg(row + i);
            }
            const int ra_val = relaxed ? (int)x0[ra_col] : 0;
            auto ev = [&](int v, int r) {
                for (int i = 0; i < D0; ++i) a[i] = x0[i];
                a[pa_col] = (float)v;
                if (relaxed) a[ra_col] = (float)r;
                FbGenReluVisit<float> vis;
                return fb_gen_forward<float>(net, net.W32, net.B32, a, y, vis);
            };
            const uint32_t low = fb_screen_sample(ev, pa_lo, W, relaxed, ra_val, ra_thr);
            const bool ce = active && low != 0xffffffffu;
            if (ce && best == ~0ull) best = ((unsigned long long)(uint32_t)s << 32) | low;
            if (__any_sync(0xffffffffu, ce)) break;
        }
        const uint32_t hi = (uint32_t)(best >> 32);
        const uint32_t hmin = __reduce_min_sync(0xffffffffu, hi);
        if (hmin != 0xffffffffu && hi == hmin) atomicMin(&ce_key[p], best);
    }
}
