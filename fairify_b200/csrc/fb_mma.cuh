// fb_mma.cuh -- activation profile on the FP64 matrix instruction (DMMA), runtime layer widths.
//
// reference: candidate_dead_nodes, utils/prune.py:168-192 (activation != 0 counted per neuron, output layer included)
// over the samples of simluate_data (:205-222); forward pass = layer_net, utils/AC-1-Model-Functions.py:16-24.
//
// mma.sync.aligned.m8n8k4.f64 computes, for each of its 8x8 outputs, c <- fma(a_k, b_k, c) for k = 0, 1, 2, 3 in that
// order -- measured on B200 (tools/ubench_dmma.cu: 3.8 million outputs with operands of widely different magnitude,
// every one bit-identical to the ascending chain, none systematically equal to any other order) -- which is exactly the
// arithmetic contract of the other profile kernels and of the oracle: one fused multiply-add per term, terms in
// ascending input order, accumulator starting at 0, bias added last.  The instruction runs on the FP64 pipe at the DFMA
// rate (37.1 TFLOP/s measured) but needs one issue slot per 256 multiply-adds instead of one per 32, and one operand
// word per 8 multiply-adds, so a layer evaluated as [samples x inputs] x [inputs x neurons] tiles is bound by the pipe
// itself rather than by instruction issue or shared-memory bandwidth.
//
// Mapping.  A warp takes MS blocks of 8 samples of one partition through the whole network on its own, and the
// activations never leave its registers.  The instruction's operand layouts (lane 4g+t):
//     A (8 samples x 4 inputs):    lane holds (sample g, input t)
//     B (4 inputs x 8 columns):    lane holds (input t, column g)
//     C/D (8 samples x 8 columns): lane holds (sample g, columns 2t and 2t+1)
// If column c of a weight tile stands for neuron 8*tile + (c even ? c/2 : 4 + c/2) -- the columns of every tile stored in
// the order 0 4 1 5 2 6 3 7 -- then lane 4g+t ends up with neurons 8*tile + t and 8*tile + 4 + t of sample g: exactly the
// two A operands (input blocks 2*tile and 2*tile+1) the next layer needs from it.  So after bias and ReLU the accumulator
// registers of layer l ARE the A operands of layer l+1; there is no shared-memory round trip, no shuffle, and the only
// shared-memory traffic is the weights (B operands, stored lane by lane so that a load is one contiguous 256-byte row).
// Inputs are padded with zero weights to a multiple of 4, neurons with zero weights and biases to a multiple of 8
// (fma(0, x, c) = c exactly, and a padded neuron's activation is relu(0) = 0).  The single output neuron is one more
// DMMA chain whose weight tile has the logit's weights in column 0.
#pragma once
#include "fb_common.cuh"

struct FbMmaNet {
    int L, N;
    int dims[FB_MAX_LAYERS + 1];
    int kb[FB_MAX_LAYERS + 1];       // layer l: input blocks = ceil(dims[l] / 4)
    int nb[FB_MAX_LAYERS + 1];       // layer l: neuron tiles = ceil(dims[l+1] / 8)   (1 for the logit)
    int woff[FB_MAX_LAYERS + 1];     // offset (doubles) of layer l in Wf
    int bpoff[FB_MAX_LAYERS + 1];    // offset of layer l in Bp and in the per-warp counters (nb[l] * 8 entries)
    int boff[FB_MAX_LAYERS + 1];     // neuron numbering of the outputs (counts)
    int maxnb;                       // widest layer in tiles (inputs included: ceil(kb[0] / 2))
    int wtotal, btotal;              // doubles in Wf / Bp
    const double *Wf;                // layer l: [cg][kb][nt][32]: column group cg = tiles 8cg .. 8cg+7 (the last group
                                     // may hold fewer), lane 4g+t of tile nt = W[4kb+t][neuron of column g], 0 outside
    const double *Bp;                // biases in column order
};

#define FB_MMA_WARPS 8
#define FB_MMA_NT 8                  // neuron tiles per column group

// neuron (within its tile) that column c stands for
__host__ __device__ inline int fb_mma_col_neuron(int c) { return (c & 1) ? 4 + (c >> 1) : (c >> 1); }

// (volatile: the operand loads and the matrix instructions keep the order they are written in -- the loads of input
//  block k+1 in front of the instructions of block k -- instead of ptxas' choice of one load right before each use)
__device__ __forceinline__ void fb_dmma(double &d0, double &d1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}
__device__ __forceinline__ double fb_lds64(uint32_t addr)
{
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ double fb_ldg64(const double *p)
{
    double v;
    asm volatile("ld.global.nc.f64 %0, [%1];" : "=d"(v) : "l"(p));
    return v;
}

// bytes of dynamic shared memory: weights + biases + per-warp sample staging rows and counters
__host__ __device__ inline size_t fb_mma_smem_bytes(const FbMmaNet &net, int MS, int warps, bool resident)
{
    return sizeof(double) * ((resident ? (size_t)net.wtotal : 0) + net.btotal + (size_t)warps * MS * net.kb[0] * 32) +
           sizeof(uint32_t) * warps * net.btotal;
}

struct FbMmaLane {
    int g, t;
    uint32_t cmask;       // lanes that hold the same columns for the other samples of the block
};

// One layer for the warp's MS x 8 samples: hout[ms][tile] = relu(hin x W + b) (the logit: no relu, `nonzero` test),
// counts of non-zero activations added to cnt (column order).  hin[ms][i][e] is input block 2i+e, hout[ms][tile][e]
// columns 2t+e of the tile = input block 2*tile+e of the next layer.  CAP = widest layer in tiles this instantiation
// can hold; all loops over tiles and input blocks are unrolled against it, with warp-uniform exits at the real sizes.
template <int MS, int CAP, bool RESIDENT, bool LOGIT>
__device__ __forceinline__ void fb_mma_layer(const double *Wl, const double *Bl, uint32_t *cnt, int KB, int NB,
                                             const double (&hin)[MS][CAP][2], double (&hout)[MS][CAP][2],
                                             const FbMmaLane &ln, int lane, const bool (&live)[MS])
{
    constexpr int NG = LOGIT ? 1 : (CAP + FB_MMA_NT - 1) / FB_MMA_NT;
    constexpr int NTC = LOGIT ? 1 : FB_MMA_NT;
#pragma unroll
    for (int G = 0; G < NG; ++G) {
        const int tile0 = G * FB_MMA_NT;
        if (tile0 >= NB) break;
        const int ntg = min(FB_MMA_NT, NB - tile0);
#pragma unroll
        for (int ms = 0; ms < MS; ++ms)
#pragma unroll
            for (int nt = 0; nt < NTC; ++nt)
                if (tile0 + nt < CAP) { hout[ms][tile0 + nt][0] = 0.0; hout[ms][tile0 + nt][1] = 0.0; }
        const double *wg = Wl + tile0 * KB * 32 + lane;                       // (all groups before this one are full)
        uint32_t wa = RESIDENT ? (uint32_t)__cvta_generic_to_shared(wg) : 0u;
        double b0[NTC], b1[NTC];
#define FB_MMA_LOADB(B)                                                                                          \
    {                                                                                                            \
        _Pragma("unroll") for (int nt = 0; nt < NTC; ++nt)                                                       \
            if (nt < ntg) B[nt] = RESIDENT ? fb_lds64(wa + nt * 256) : fb_ldg64(wg + nt * 32);                   \
        wa += ntg * 256; wg += ntg * 32;                                                                         \
    }
#define FB_MMA_GO(B, KBI)                                                                                        \
    {                                                                                                            \
        _Pragma("unroll") for (int nt = 0; nt < NTC; ++nt)                                                       \
            if (tile0 + nt < CAP && nt < ntg) {                                                                  \
                _Pragma("unroll") for (int ms = 0; ms < MS; ++ms)                                                \
                    fb_dmma(hout[ms][tile0 + nt][0], hout[ms][tile0 + nt][1], hin[ms][(KBI) >> 1][(KBI) & 1], B[nt]); \
            }                                                                                                    \
    }
        FB_MMA_LOADB(b0);
#pragma unroll
        for (int kb = 0; kb < 2 * CAP; kb += 2) {
            if (kb >= KB) break;
            if (kb + 1 < KB) FB_MMA_LOADB(b1);
            FB_MMA_GO(b0, kb);
            if (kb + 1 < KB) {
                if (kb + 2 < KB) FB_MMA_LOADB(b0);
                FB_MMA_GO(b1, kb + 1);
            }
        }
#undef FB_MMA_LOADB
#undef FB_MMA_GO
        // bias, sign, counts, ReLU in place.  Lane (g, t) keeps the counts of tile g of this group, columns 2t and 2t+1.
        uint32_t my0 = 0, my1 = 0;
#pragma unroll
        for (int nt = 0; nt < NTC; ++nt) {
            if (tile0 + nt < CAP && nt < ntg) {
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
        if (ln.g < ntg) {
            uint2 *cp = reinterpret_cast<uint2 *>(cnt + (tile0 + ln.g) * 8 + 2 * ln.t);
            uint2 v = *cp;
            v.x += my0; v.y += my1;
            *cp = v;
        }
    }
}

template <int MS, int CAP, bool PHILOX, bool RESIDENT>
__global__ void __launch_bounds__(FB_MMA_WARPS * 32)
fb_profile_mma_kernel(const FbMmaNet net, int64_t n, const int32_t *__restrict__ boxes,
                      const int32_t *__restrict__ samples, uint64_t seed, int64_t first_index, int nsamp, int chunks,
                      int chunk_len, uint32_t *__restrict__ counts)
{
    extern __shared__ __align__(16) unsigned char fb_smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;
    FbMmaLane ln;
    ln.g = lane >> 2; ln.t = lane & 3; ln.cmask = 0x11111111u << ln.t;
    const int L = net.L, N = net.N, D0 = net.dims[0];
    double *sW = reinterpret_cast<double *>(fb_smem);
    const int wres = RESIDENT ? net.wtotal : 0;
    double *sB = sW + wres;
    double *sStage = sB + net.btotal + warp * (MS * net.kb[0] * 32);         // this warp's sample staging rows
    uint32_t *sCnt = reinterpret_cast<uint32_t *>(sB + net.btotal + nwarp * (MS * net.kb[0] * 32)) + warp * net.btotal;
    for (int i = tid; i < wres; i += blockDim.x) sW[i] = __ldg(net.Wf + i);
    for (int i = tid; i < net.btotal; i += blockDim.x) sB[i] = __ldg(net.Bp + i);
    for (int k = lane; k < net.btotal; k += 32) sCnt[k] = 0;
    __syncthreads();
    const double *Wbase = RESIDENT ? sW : net.Wf;

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
        const int KB0 = net.kb[0];
        for (int sb0 = s0; sb0 < s1; sb0 += MS * 8) {
            double ha[MS][CAP][2], hb[MS][CAP][2];
            bool live[MS];
            // ---- this step's samples: lanes 0 .. 8*MS-1 draw (or fetch) one sample each and leave its features in the
            // warp's staging rows in operand order -- (sample 8ms+gg, feature i) at ((ms*KB0 + i/4)*32 + 4gg + i%4) --
            // from where every lane picks up its A operands: lane (g, t) holds feature 4kb + t of sample sb0 + 8ms + g
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
                        for (int i = 0; i < D0; ++i) row[(i >> 2) * 32 + (i & 3)] = (double)__ldg(src + i);
                    }
                    for (int i = D0; i < KB0 * 4; ++i) row[(i >> 2) * 32 + (i & 3)] = 0.0;
                } else {
                    for (int i = 0; i < KB0 * 4; ++i) row[(i >> 2) * 32 + (i & 3)] = 0.0;
                }
            }
            __syncwarp();
#pragma unroll
            for (int ms = 0; ms < MS; ++ms) {
                live[ms] = sb0 + ms * 8 + ln.g < s1;
#pragma unroll
                for (int kb = 0; kb < 2 * CAP; ++kb) {
                    if (kb >= KB0) break;
                    ha[ms][kb >> 1][kb & 1] = sStage[(ms * KB0 + kb) * 32 + lane];
                }
            }
            __syncwarp();
            // ---- layers in pairs (the two register arrays swap roles)
#pragma unroll 1
            for (int l = 0;; l += 2) {
                if (l + 1 < L) fb_mma_layer<MS, CAP, RESIDENT, false>(Wbase + net.woff[l], sB + net.bpoff[l], sCnt + net.bpoff[l], net.kb[l], net.nb[l], ha, hb, ln, lane, live);
                else { fb_mma_layer<MS, CAP, RESIDENT, true>(Wbase + net.woff[l], sB + net.bpoff[l], sCnt + net.bpoff[l], net.kb[l], 1, ha, hb, ln, lane, live); break; }
                if (l + 2 < L) fb_mma_layer<MS, CAP, RESIDENT, false>(Wbase + net.woff[l + 1], sB + net.bpoff[l + 1], sCnt + net.bpoff[l + 1], net.kb[l + 1], net.nb[l + 1], hb, ha, ln, lane, live);
                else { fb_mma_layer<MS, CAP, RESIDENT, true>(Wbase + net.woff[l + 1], sB + net.bpoff[l + 1], sCnt + net.bpoff[l + 1], net.kb[l + 1], 1, hb, ha, ln, lane, live); break; }
            }
        }
        __syncwarp();
        for (int l = 0; l < L; ++l) {
            for (int j = lane; j < net.dims[l + 1]; j += 32) {
                // neuron j sits in column (j % 8 < 4 ? 2 * (j % 8) : 2 * (j % 8 - 4) + 1) of tile j / 8
                const int cj = (j & ~7) + ((j & 4) ? 2 * (j & 3) + 1 : 2 * (j & 3));
                const uint32_t v = sCnt[net.bpoff[l] + cj];
                const int64_t k = p * N + net.boff[l] + j;
                if (chunks == 1) counts[k] = v;
                else if (v) atomicAdd(&counts[k], v);
                sCnt[net.bpoff[l] + cj] = 0;
            }
        }
        __syncwarp();
    }
}
