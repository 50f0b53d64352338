// fb_tiled.cuh -- activation-profile kernel for networks of any shape (runtime widths).
//
// reference: candidate_dead_nodes, utils/prune.py:168-192 over the samples of simluate_data (:205-222).
//
// One warp per (partition, sample chunk); each lane carries K samples per step.  All data a lane touches
// besides the weights is private to it and lives in shared memory in lane-major columns (no bank
// conflicts, no synchronisation): the activations of the current and the next layer, and one 8-bit
// non-zero counter per neuron.  A layer is evaluated in blocks of JB output neurons (16, 8, 4, 2, 1 -- picked
// per block, so no padding work): the JB x K accumulators stay in registers while a rolled loop walks the
// inputs i = 0, 1, ...; each iteration loads the lane's K activations x_i and the weight row slice
// w[i][j0 .. j0+JB) with 128-bit shared-memory loads (every lane reads the same address: one wavefront) and
// issues JB*K fused multiply-adds.  Per output neuron the terms are added for i ascending with one FMA
// each and the bias last: bit-identical to fb_spec_forward and to the oracle.
//
// Why shared memory and rolled loops rather than the constant bank and full unrolling of fb_spec.cuh: the
// compiler places each uniform constant load (LDCU) directly in front of the DFMA that consumes it, so every
// weight costs an exposed load latency; ordinary LDS are hoisted to the top of the small loop body, where one
// wait covers JB*K FMAs.  The loops also keep the code a few KB instead of tens of KB.
#pragma once
#include "fb_common.cuh"

#define FB_TILED_MAX_STEPS(K) (255 / (K))     // 8-bit counters, K increments per step at most

struct TiledNet {
    int L, N, P, Ppad, d;
    int dims[FB_MAX_LAYERS + 1];
    int ld[FB_MAX_LAYERS];            // padded row length (even) of layer l's weights
    int woff[FB_MAX_LAYERS + 1];      // offsets into the padded weight array
    int boff[FB_MAX_LAYERS + 1];
    int wA, wB;                       // rows of the two activation buffers (even / odd layer inputs)
    const double *Wpad, *B64;         // device: padded weights, biases
};

template <int JB, int K, int THREADS>
__device__ __forceinline__ void fb_tile_block(const double *__restrict__ w, int ld, int IN, const double *actin,
                                              double (&acc)[JB][K])
{
#pragma unroll
    for (int jj = 0; jj < JB; ++jj)
#pragma unroll
        for (int k = 0; k < K; ++k) acc[jj][k] = 0.0;
#pragma unroll 2
    for (int i = 0; i < IN; ++i) {
        double x[K];
#pragma unroll
        for (int k = 0; k < K; ++k) x[k] = actin[(i * K + k) * THREADS];
        const double *wr = w + i * ld;
        if constexpr (JB >= 2) {
#pragma unroll
            for (int jj = 0; jj < JB; jj += 2) {
                const double2 ww = *reinterpret_cast<const double2 *>(wr + jj);
#pragma unroll
                for (int k = 0; k < K; ++k) {
                    acc[jj][k] = __fma_rn(ww.x, x[k], acc[jj][k]);
                    acc[jj + 1][k] = __fma_rn(ww.y, x[k], acc[jj + 1][k]);
                }
            }
        } else {
            const double ww = wr[0];
#pragma unroll
            for (int k = 0; k < K; ++k) acc[0][k] = __fma_rn(ww, x[k], acc[0][k]);
        }
    }
}

// bias, ReLU, non-zero counting and hand-over to the next layer for one block of JB neurons
template <int JB, int K, int THREADS>
__device__ __forceinline__ void fb_tile_epilogue(const double (&acc)[JB][K], const double *__restrict__ bias, bool last,
                                                 const bool (&active)[K], double *actout, uint8_t *cnt)
{
#pragma unroll
    for (int jj = 0; jj < JB; ++jj) {
        const double bj = bias[jj];
        uint32_t c = 0;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const double pre = __dadd_rn(acc[jj][k], bj);
            const bool pos = fb_pos(pre);
            const bool nz = last ? fb_nonzero(pre) : pos;
            c += (nz && active[k]) ? 1u : 0u;
            if (!last) actout[(jj * K + k) * THREADS] = pos ? pre : 0.0;
        }
        cnt[jj * THREADS] = (uint8_t)(cnt[jj * THREADS] + c);
    }
}

template <int JB, int K, int THREADS>
__device__ __forceinline__ void fb_tile_run(const double *__restrict__ w, int ld, int IN, const double *actin,
                                            const double *__restrict__ bias, bool last, const bool (&active)[K],
                                            double *actout, uint8_t *cnt)
{
    double acc[JB][K];
    fb_tile_block<JB, K, THREADS>(w, ld, IN, actin, acc);
    fb_tile_epilogue<JB, K, THREADS>(acc, bias, last, active, actout, cnt);
}

template <int K, int THREADS, bool PHILOX>
__global__ void __launch_bounds__(THREADS)
fb_profile_tiled_kernel(const TiledNet net, int64_t n, const int32_t *__restrict__ boxes,
                        const int32_t *__restrict__ samples, uint64_t seed, int64_t first_index, int nsamp, int chunks,
                        int chunk_len, uint32_t *__restrict__ counts)
{
    extern __shared__ __align__(16) unsigned char fb_smem[];
    double *sW = reinterpret_cast<double *>(fb_smem);
    double *sB = sW + net.Ppad;
    double *sA = sB + ((net.N + 1) & ~1);
    double *sBb = sA + (size_t)net.wA * K * THREADS;
    uint8_t *sCnt = reinterpret_cast<uint8_t *>(sBb + (size_t)net.wB * K * THREADS);
    int2 *sBox = reinterpret_cast<int2 *>(sCnt + (((size_t)net.N * THREADS + 15) & ~(size_t)15));
    for (int k = threadIdx.x; k < net.Ppad; k += THREADS) sW[k] = net.Wpad[k];
    for (int k = threadIdx.x; k < net.N; k += THREADS) sB[k] = net.B64[k];
    __syncthreads();

    const int tid = threadIdx.x, lane = tid & 31, wib = tid >> 5;
    const int D0 = net.d, N = net.N;
    double *actA = sA + tid, *actB = sBb + tid;
    uint8_t *cnt = sCnt + tid;
    int2 *box = sBox + wib * D0;
    const int64_t nwarps = (int64_t)gridDim.x * (THREADS / 32);
    const int64_t gwarp = (int64_t)blockIdx.x * (THREADS / 32) + wib;
    const int64_t units = n * chunks;
    for (int64_t u = gwarp; u < units; u += nwarps) {
        const int64_t p = u / chunks;
        const int c = (int)(u - p * chunks);
        const int s0 = c * chunk_len;
        const int s1 = min(nsamp, s0 + chunk_len);
        const int64_t pidx = first_index + p;
        if (PHILOX) {
            __syncwarp();
            for (int i = lane; i < D0; i += 32) {
                const int2 b = __ldg(reinterpret_cast<const int2 *>(boxes + p * (2 * D0)) + i);
                box[i] = make_int2(b.x, b.y - b.x + 1);          // (lo, range)
            }
            __syncwarp();
        }
        for (int j = 0; j < N; ++j) cnt[j * THREADS] = 0;
        for (int sb = s0; sb < s1; sb += 32 * K) {
            bool active[K];
#pragma unroll
            for (int k = 0; k < K; ++k) active[k] = (sb + 32 * k + lane) < s1;
            // layer-0 inputs -> buffer A
            if (PHILOX) {
                for (int blk = 0; blk * 4 < D0; ++blk) {
                    uint32_t r[K][4];
#pragma unroll
                    for (int k = 0; k < K; ++k)
                        fb_philox4x32_10((uint32_t)(sb + 32 * k + lane), (uint32_t)pidx, (uint32_t)((uint64_t)pidx >> 32),
                                         (uint32_t)blk, (uint32_t)seed, (uint32_t)(seed >> 32), r[k]);
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int i = blk * 4 + q;
                        if (i < D0) {
                            const int2 lr = box[i];
#pragma unroll
                            for (int k = 0; k < K; ++k)
                                actA[(i * K + k) * THREADS] = (double)fb_map_range(r[k][q], lr.x, (uint32_t)lr.y);
                        }
                    }
                }
            } else {
#pragma unroll
                for (int k = 0; k < K; ++k) {
                    const int s = min(sb + 32 * k + lane, s1 - 1);
                    const int32_t *row = samples + ((int64_t)p * nsamp + s) * D0;
                    for (int i = 0; i < D0; ++i) actA[(i * K + k) * THREADS] = (double)__ldg(row + i);
                }
            }
            // layers
            double *in = actA, *out = actB;
            for (int l = 0; l < net.L; ++l) {
                const int IN = net.dims[l], OUT = net.dims[l + 1], ld = net.ld[l], bo = net.boff[l];
                const double *wl = sW + net.woff[l];
                const bool last = (l == net.L - 1);
                int j0 = 0;
                while (j0 < OUT) {
                    const int rem = OUT - j0;
                    const double *w = wl + j0, *bias = sB + bo + j0;
                    double *ao = out + (size_t)j0 * K * THREADS;
                    uint8_t *cn = cnt + (size_t)(bo + j0) * THREADS;
                    if (rem >= 16)     { fb_tile_run<16, K, THREADS>(w, ld, IN, in, bias, last, active, ao, cn); j0 += 16; }
                    else if (rem >= 8) { fb_tile_run<8, K, THREADS>(w, ld, IN, in, bias, last, active, ao, cn); j0 += 8; }
                    else if (rem >= 4) { fb_tile_run<4, K, THREADS>(w, ld, IN, in, bias, last, active, ao, cn); j0 += 4; }
                    else if (rem >= 2) { fb_tile_run<2, K, THREADS>(w, ld, IN, in, bias, last, active, ao, cn); j0 += 2; }
                    else               { fb_tile_run<1, K, THREADS>(w, ld, IN, in, bias, last, active, ao, cn); j0 += 1; }
                }
                double *t = in; in = out; out = t;
            }
        }
        // per-lane counters -> per-partition counts
        uint32_t mine = 0;
        for (int j = 0; j < N; ++j) {
            const uint32_t tot = __reduce_add_sync(0xffffffffu, (uint32_t)cnt[j * THREADS]);
            if ((j & 31) == lane) mine = tot;
            if ((j & 31) == 31 || j == N - 1) {
                const int jj = (j & ~31) + lane;
                if (jj <= j) {
                    if (chunks == 1) counts[p * N + jj] = mine;
                    else atomicAdd(&counts[p * N + jj], mine);
                }
            }
        }
    }
}
