// fb_mma_spec.cuh -- the DMMA activation-profile kernel (fb_mma.cuh) with the layer widths as template constants.
//
// Same mapping, same operand layouts, same arithmetic as fb_profile_mma_kernel; what the compile-time shape buys is a
// straight-line instruction stream: every input block and neuron tile is an unrolled step, so a layer is one weight
// load and one matrix instruction per 256 multiply-adds with no loop counters, bounds tests or re-convergence points
// in between, and the register arrays that carry the activations have exactly the size of each layer.
#pragma once
#include "fb_mma.cuh"

template <class S> __host__ __device__ constexpr int fb_mma_kb(int l) { return (S::dim(l) + 3) / 4; }
template <class S> __host__ __device__ constexpr int fb_mma_nb(int l) { return (S::dim(l + 1) + 7) / 8; }
template <class S> __host__ __device__ constexpr int fb_mma_woff(int l)
{
    int o = 0;
    for (int k = 0; k < l; ++k) o += fb_mma_kb<S>(k) * fb_mma_nb<S>(k) * 32;
    return o;
}
template <class S> __host__ __device__ constexpr int fb_mma_bpoff(int l)
{
    int o = 0;
    for (int k = 0; k < l; ++k) o += fb_mma_nb<S>(k) * 8;
    return o;
}

// Layer l and everything after it.  hin[ms][i][e] = input block 2i+e of the layer (see fb_mma.cuh).
// (ptxas re-sorts the unrolled matrix instructions of a column group tile by tile -- one accumulator live at a time, each
//  tile a chain of dependent DMMAs -- whatever order they are written in.  Forcing all tiles to advance together (a
//  never-taken run-time exit per input block, which pins the written order) was measured on B200: it costs 20
//  registers, which for the 64-wide networks is the second resident CTA, and loses 6-9 points; with 16 warps per SM the
//  serial chains are covered by the other warps.)
template <class S, int l, int MS, int NBIN>
__device__ __forceinline__ void fb_mma_spec_layers(const double *sW, const double *sB, uint32_t *sCnt,
                                                   const double (&hin)[MS][NBIN][2], const FbMmaLane &ln, int lane,
                                                   const bool (&live)[MS])
{
    constexpr int KB = fb_mma_kb<S>(l), NB = fb_mma_nb<S>(l);
    constexpr bool LOGIT = (l == S::NL - 1);
    static_assert(2 * NBIN >= KB, "input operand array too small");
    const double *Wl = sW + fb_mma_woff<S>(l);
    const double *Bl = sB + fb_mma_bpoff<S>(l);
    uint32_t *cnt = sCnt + fb_mma_bpoff<S>(l);
    double hout[MS][NB][2];
#pragma unroll
    for (int ms = 0; ms < MS; ++ms)
#pragma unroll
        for (int nt = 0; nt < NB; ++nt) { hout[ms][nt][0] = 0.0; hout[ms][nt][1] = 0.0; }
#pragma unroll
    for (int tile0 = 0; tile0 < NB; tile0 += FB_MMA_NT) {
        constexpr int dummy = 0; (void)dummy;
        const int NTG = (NB - tile0 < FB_MMA_NT) ? NB - tile0 : FB_MMA_NT;         // compile-time after unrolling
        uint32_t wa = (uint32_t)__cvta_generic_to_shared(Wl + tile0 * KB * 32 + lane);
        double b0[FB_MMA_NT], b1[FB_MMA_NT];
#pragma unroll
        for (int nt = 0; nt < FB_MMA_NT; ++nt)
            if (nt < NTG) b0[nt] = fb_lds64(wa + nt * 256);
#pragma unroll
        for (int kb = 0; kb < KB; kb += 2) {
            if (kb + 1 < KB) {
#pragma unroll
                for (int nt = 0; nt < FB_MMA_NT; ++nt)
                    if (nt < NTG) b1[nt] = fb_lds64(wa + ((kb + 1) * NTG + nt) * 256);
            }
#pragma unroll
            for (int nt = 0; nt < FB_MMA_NT; ++nt)
                if (nt < NTG) {
#pragma unroll
                    for (int ms = 0; ms < MS; ++ms)
                        fb_dmma(hout[ms][tile0 + nt][0], hout[ms][tile0 + nt][1], hin[ms][kb >> 1][kb & 1], b0[nt]);
                }
            if (kb + 1 < KB) {
                if (kb + 2 < KB) {
#pragma unroll
                    for (int nt = 0; nt < FB_MMA_NT; ++nt)
                        if (nt < NTG) b0[nt] = fb_lds64(wa + ((kb + 2) * NTG + nt) * 256);
                }
#pragma unroll
                for (int nt = 0; nt < FB_MMA_NT; ++nt)
                    if (nt < NTG) {
#pragma unroll
                        for (int ms = 0; ms < MS; ++ms)
                            fb_dmma(hout[ms][tile0 + nt][0], hout[ms][tile0 + nt][1], hin[ms][(kb + 1) >> 1][(kb + 1) & 1], b1[nt]);
                    }
            }
        }
        // bias, sign, counts, ReLU in place.  Lane (g, t) keeps the counts of tile g of this group, columns 2t and 2t+1.
        uint32_t my0 = 0, my1 = 0;
#pragma unroll
        for (int nt = 0; nt < FB_MMA_NT; ++nt) {
            if (nt < NTG) {
                const double2 bj = *reinterpret_cast<const double2 *>(Bl + (tile0 + nt) * 8 + 2 * ln.t);
                uint32_t c0 = 0, c1 = 0;
#pragma unroll
                for (int ms = 0; ms < MS; ++ms) {
                    const double p0 = __dadd_rn(hout[ms][tile0 + nt][0], bj.x), p1 = __dadd_rn(hout[ms][tile0 + nt][1], bj.y);
                    const bool q0 = LOGIT ? fb_nonzero(p0) : fb_pos(p0), q1 = LOGIT ? false : fb_pos(p1);
                    const uint32_t m0 = __ballot_sync(0xffffffffu, q0 && live[ms]);
                    c0 += __popc(m0 & ln.cmask);
                    if (!LOGIT) {
                        const uint32_t m1 = __ballot_sync(0xffffffffu, q1 && live[ms]);
                        c1 += __popc(m1 & ln.cmask);
                        hout[ms][tile0 + nt][0] = q0 ? p0 : 0.0; hout[ms][tile0 + nt][1] = q1 ? p1 : 0.0;
                    }
                }
                if (nt == ln.g) { my0 = c0; my1 = c1; }
            }
        }
        if (ln.g < NTG) {
            uint2 *cp = reinterpret_cast<uint2 *>(cnt + (tile0 + ln.g) * 8 + 2 * ln.t);
            uint2 v = *cp;
            v.x += my0; v.y += my1;
            *cp = v;
        }
    }
    if constexpr (!LOGIT) fb_mma_spec_layers<S, l + 1, MS, NB>(sW, sB, sCnt, hout, ln, lane, live);
}

// widest layer of S in neuron tiles
template <class S> __host__ __device__ constexpr int fb_mma_maxnb()
{
    int r = 0;
    for (int l = 0; l < S::NL; ++l) r = fb_mma_nb<S>(l) > r ? fb_mma_nb<S>(l) : r;
    return r;
}

// two CTAs per SM (registers capped at 128) for one sample block of a network up to 64 wide: the matrix-instruction
// stream of a warp has gaps (bias / sign / count epilogues, sample generation) that only other resident warps fill
template <class S, int MS, bool PHILOX>
__global__ void __launch_bounds__(FB_MMA_WARPS * 32, (MS == 1 && fb_mma_maxnb<S>() <= 8) ? 2 : 1)
fb_profile_mma_spec_kernel(const FbMmaNet net, int64_t n, const int32_t *__restrict__ boxes,
                           const int32_t *__restrict__ samples, uint64_t seed, int64_t first_index, int nsamp, int chunks,
                           int chunk_len, uint32_t *__restrict__ counts)
{
    extern __shared__ __align__(16) unsigned char fb_smem[];
    constexpr int D0 = S::D0, N = S::N, L = S::NL, KB0 = fb_mma_kb<S>(0), NBIN0 = (KB0 + 1) / 2;
    constexpr int WTOTAL = fb_mma_woff<S>(L), BTOTAL = fb_mma_bpoff<S>(L);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;
    FbMmaLane ln;
    ln.g = lane >> 2; ln.t = lane & 3; ln.cmask = 0x11111111u << ln.t;
    double *sW = reinterpret_cast<double *>(fb_smem);
    double *sB = sW + WTOTAL;
    double *sStage = sB + BTOTAL + warp * (MS * KB0 * 32);                  // this warp's sample staging rows
    uint32_t *sCnt = reinterpret_cast<uint32_t *>(sB + BTOTAL + nwarp * (MS * KB0 * 32)) + warp * BTOTAL;
    for (int i = tid; i < WTOTAL; i += blockDim.x) sW[i] = __ldg(net.Wf + i);
    for (int i = tid; i < BTOTAL; i += blockDim.x) sB[i] = __ldg(net.Bp + i);
    for (int k = lane; k < BTOTAL; k += 32) sCnt[k] = 0;
    __syncthreads();

    const int64_t nwarps = (int64_t)gridDim.x * nwarp;
    const int64_t gwarp = (int64_t)blockIdx.x * nwarp + warp;
    const int64_t units = n * chunks;
    for (int64_t u = gwarp; u < units; u += nwarps) {
        const int64_t p = u / chunks;
        const int c = (int)(u - p * chunks);
        const int s0 = c * chunk_len, s1 = min(nsamp, s0 + chunk_len);
        const int32_t *box = boxes + p * (2 * D0);
        const int64_t pidx = first_index + p;
        const int grp = PHILOX ? fb_draw_group(fb_box_widest(box, D0)) : 1;
        for (int sb0 = s0; sb0 < s1; sb0 += MS * 8) {
            // ---- this step's samples: lanes 0 .. 8*MS-1 draw (or fetch) one sample each and leave its features in the
            // warp's staging rows in operand order, from where every lane picks up its A operands
            if (lane < MS * 8) {
                const int s = sb0 + lane;
                double *row = sStage + ((lane >> 3) * KB0) * 32 + 4 * (lane & 7);
                if (s < s1) {
                    if (PHILOX) {
                        const int nwords = (D0 + grp - 1) / grp;
                        for (int blk = 0; blk * 4 < nwords; ++blk) {
                            uint32_t r[4];
                            fb_philox4x32_10((uint32_t)s, (uint32_t)pidx, (uint32_t)((uint64_t)pidx >> 32), (uint32_t)blk,
                                             (uint32_t)seed, (uint32_t)(seed >> 32), r);
#pragma unroll
                            for (int k = 0; k < 4; ++k) {
                                uint32_t word = r[k];
#pragma unroll
                                for (int q = 0; q < 4; ++q) {
                                    const int i = (blk * 4 + k) * grp + q;
                                    if (q < grp && i < D0) {
                                        const int lo = __ldg(box + 2 * i), hi = __ldg(box + 2 * i + 1);
                                        row[(i >> 2) * 32 + (i & 3)] = (double)fb_draw_next(word, lo, (uint32_t)(hi - lo + 1));
                                    }
                                }
                            }
                        }
                    } else {
                        const int32_t *src = samples + ((int64_t)p * nsamp + s) * D0;
#pragma unroll
                        for (int i = 0; i < D0; ++i) row[(i >> 2) * 32 + (i & 3)] = (double)__ldg(src + i);
                    }
#pragma unroll
                    for (int i = D0; i < KB0 * 4; ++i) row[(i >> 2) * 32 + (i & 3)] = 0.0;
                } else {
#pragma unroll
                    for (int i = 0; i < KB0 * 4; ++i) row[(i >> 2) * 32 + (i & 3)] = 0.0;
                }
            }
            __syncwarp();
            double ha[MS][NBIN0][2];
            bool live[MS];
#pragma unroll
            for (int ms = 0; ms < MS; ++ms) {
                live[ms] = sb0 + ms * 8 + ln.g < s1;
#pragma unroll
                for (int kb = 0; kb < 2 * NBIN0; ++kb) ha[ms][kb >> 1][kb & 1] = kb < KB0 ? sStage[(ms * KB0 + kb) * 32 + lane] : 0.0;
            }
            __syncwarp();
            fb_mma_spec_layers<S, 0, MS, NBIN0>(sW, sB, sCnt, ha, ln, lane, live);
        }
        __syncwarp();
#pragma unroll
        for (int l = 0; l < L; ++l) {
            for (int j = lane; j < S::dim(l + 1); j += 32) {
                const int cj = (j & ~7) + ((j & 4) ? 2 * (j & 3) + 1 : 2 * (j & 3));
                const uint32_t v = sCnt[fb_mma_bpoff<S>(l) + cj];
                const int64_t k = p * N + S::boff(l) + j;
                if (chunks == 1) counts[k] = v;
                else if (v) atomicAdd(&counts[k], v);
                sCnt[fb_mma_bpoff<S>(l) + cj] = 0;
            }
        }
        __syncwarp();
    }
}
