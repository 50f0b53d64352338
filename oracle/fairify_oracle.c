/*
 * fairify_oracle.c -- CPU restatement of Fairify's pre-solver stage.
 *
 * TEST INFRASTRUCTURE ONLY.  This file is the checker the CUDA path is compared against; it is never
 * on the product path.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load it.
 *
 * Every function cites the reference code it restates (paths relative to the Fairify checkout).
 * Parity status: PINNED against the reference itself.  tests/golden/make_golden.py imports the
 * reference's utils/prune.py, utils/input_partition.py and utils/<M>-Model-Functions.py unmodified
 * (tensorflow / aif360 stubbed) in the build container, runs them on the real model weights, and the
 * committed fixtures are checked against this file by tests/test_oracle_golden.py.  Two boundaries stay
 * unpinned because the code behind them is not in the reference tree: the summation order of numpy's
 * BLAS dgemv (layer_net / net) and TensorFlow 2.5's float32 Dense kernels (pair screen).  For both this
 * file fixes the order to "k ascending, one fused multiply-add per term, bias added last"; only the
 * sign of a pre-activation / logit is ever consumed, so the two can differ only for values within a
 * few ulp of zero (none in the fixtures).
 *
 * Build: make -C oracle   (gcc -O2 -mfma -ffp-contract=off -fopenmp)
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

#define FO_MAX_LAYERS 16

typedef struct {
    int L;                       /* weight layers */
    const int32_t *dims;         /* L+1 entries: d, h1, ..., 1 */
    const double *W;             /* layer-major, each [in][out] row-major (Keras kernel layout) */
    const double *B;             /* layer-major biases */
    int64_t woff[FO_MAX_LAYERS + 1];
    int64_t boff[FO_MAX_LAYERS + 1];
    int N;                       /* total neurons, output neuron included */
    int maxw;
} fo_net;

static int fo_net_init(fo_net *m, int L, const int32_t *dims, const double *W, const double *B)
{
    if (L < 1 || L > FO_MAX_LAYERS) return -1;
    m->L = L; m->dims = dims; m->W = W; m->B = B;
    int64_t wo = 0, bo = 0; int maxw = dims[0];
    for (int l = 0; l < L; ++l) {
        m->woff[l] = wo; m->boff[l] = bo;
        wo += (int64_t)dims[l] * dims[l + 1];
        bo += dims[l + 1];
        if (dims[l + 1] > maxw) maxw = dims[l + 1];
    }
    m->woff[L] = wo; m->boff[L] = bo; m->N = (int)bo; m->maxw = maxw;
    return 0;
}

void fo_set_threads(int n)
{
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

int fo_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* ------------------------------------------------------------------------------------------------
 * a1 / a2: partition grid.  utils/input_partition.py:17-46 (partition) splits an attribute only when
 * hi-lo+1 > T, into chunks [lo+k*T, min(lo+k*T+T-1, hi)]; :48-76 (partitioned_ranges) takes the
 * cartesian product over the split attributes in range_dict KEY order, last key fastest
 * (itertools.product), un-split attributes keep their full range.
 * key_order[k] = column index of the k-th range_dict key.  Boxes are written in COLUMN order.
 * ---------------------------------------------------------------------------------------------- */
int64_t fo_grid_count(int d, const int64_t *lo, const int64_t *hi, int64_t T, int64_t *nchunks)
{
    int64_t n = 1;
    for (int c = 0; c < d; ++c) {
        int64_t size = hi[c] - lo[c] + 1;
        int64_t k = (size <= T) ? 1 : (size + T - 1) / T;
        nchunks[c] = k;
        n *= k;
    }
    return n;
}

void fo_grid_boxes(int d, const int64_t *lo, const int64_t *hi, int64_t T, const int32_t *key_order,
                   int64_t first, int64_t count, int32_t *boxes)
{
    int64_t nch[64];
    fo_grid_count(d, lo, hi, T, nch);
    for (int64_t p = 0; p < count; ++p) {
        int64_t idx = first + p;
        int32_t *box = boxes + p * d * 2;
        for (int k = d - 1; k >= 0; --k) {           /* last key is the fastest digit */
            int c = key_order[k];
            int64_t digit = idx % nch[c];
            idx /= nch[c];
            if (nch[c] == 1) { box[2 * c] = (int32_t)lo[c]; box[2 * c + 1] = (int32_t)hi[c]; }
            else {
                int64_t a = lo[c] + digit * T, b = a + T - 1;
                if (b > hi[c]) b = hi[c];
                box[2 * c] = (int32_t)a; box[2 * c + 1] = (int32_t)b;
            }
        }
    }
}

/* Point -> partition: canonical index of the box containing each row of X (float64 [m][d], column order),
 * or -1 when no box does (outside the domain, or between two integer chunks).  Restates the membership test
 * of find_partition_result_for_point (src/AC/Verify-AC-experiment-new2.py:587-611: lower <= x <= upper for
 * every attribute) and of p_list_density (utils/input_partition.py:198-212), with the linear scan over all
 * partitions replaced by the mixed-radix index arithmetic of the grid. */
void fo_grid_locate(int d, const int64_t *lo, const int64_t *hi, int64_t T, const int32_t *key_order, int64_t m,
                    const double *X, int64_t *index)
{
    int64_t nch[64];
    fo_grid_count(d, lo, hi, T, nch);
    for (int64_t r = 0; r < m; ++r) {
        int64_t idx = 0; int ok = 1;
        for (int k = 0; k < d && ok; ++k) {               /* first key is the slowest digit */
            int c = key_order[k];
            double x = X[r * d + c];
            if (!(x >= (double)lo[c] && x <= (double)hi[c])) { ok = 0; break; }
            int64_t digit = 0;
            if (nch[c] > 1) {
                digit = (int64_t)floor((x - (double)lo[c]) / (double)T);
                int64_t top = lo[c] + digit * T + T - 1; if (top > hi[c]) top = hi[c];
                if (x > (double)top) { ok = 0; break; }
            }
            idx = idx * nch[c] + digit;
        }
        index[r] = ok ? idx : -1;
    }
}

/* ------------------------------------------------------------------------------------------------
 * a6: neuron_bounds, utils/prune.py:105-164, float64 semantics of the reference's pinned numpy
 * (weights cast to float64 first).  Per neuron: min_t = max_t = 0; for i ascending the product is
 * rounded, then added (no fused multiply-add); negative weights swap lb/ub (:139-144); bias added
 * last (:146-147); post-ReLU bound is 0 if v < 0 else v for every layer (:151-152).
 * Outputs are [n][N] (N = all neurons, layer-major).
 * ---------------------------------------------------------------------------------------------- */
static void ibp_one(const fo_net *m, const int32_t *box, double *scratch /* 4*maxw */,
                    double *ws_lb, double *ws_ub, double *pl_lb, double *pl_ub)
{
    double *lb = scratch, *ub = lb + m->maxw, *nlb = ub + m->maxw, *nub = nlb + m->maxw;
    int d = m->dims[0];
    for (int i = 0; i < d; ++i) { lb[i] = (double)box[2 * i]; ub[i] = (double)box[2 * i + 1]; }
    for (int l = 0; l < m->L; ++l) {
        int in = m->dims[l], out = m->dims[l + 1];
        const double *w = m->W + m->woff[l], *b = m->B + m->boff[l];
        for (int j = 0; j < out; ++j) {
            double mn = 0.0, mx = 0.0;
            for (int i = 0; i < in; ++i) {
                double wt = w[(int64_t)i * out + j];
                if (wt < 0) { mn += wt * ub[i]; mx += wt * lb[i]; }
                else        { mn += wt * lb[i]; mx += wt * ub[i]; }
            }
            double mnw = mn + b[j], mxw = mx + b[j];
            int64_t o = m->boff[l] + j;
            ws_lb[o] = mnw; ws_ub[o] = mxw;
            nlb[j] = mnw < 0 ? 0.0 : mnw; nub[j] = mxw < 0 ? 0.0 : mxw;
            pl_lb[o] = nlb[j]; pl_ub[o] = nub[j];
        }
        memcpy(lb, nlb, sizeof(double) * out); memcpy(ub, nub, sizeof(double) * out);
    }
}

void fo_ibp(int L, const int32_t *dims, const double *W, const double *B, int64_t n,
            const int32_t *boxes, double *ws_lb, double *ws_ub, double *pl_lb, double *pl_ub)
{
    fo_net m; if (fo_net_init(&m, L, dims, W, B)) return;
    int d = dims[0];
#pragma omp parallel
    {
        double *scratch = (double *)malloc(sizeof(double) * m.maxw * 4);
#pragma omp for schedule(static)
        for (int64_t p = 0; p < n; ++p)
            ibp_one(&m, boxes + p * d * 2, scratch, ws_lb + p * m.N, ws_ub + p * m.N,
                    pl_lb + p * m.N, pl_ub + p * m.N);
        free(scratch);
    }
}

/* Directed rounding without touching the FP environment: round-to-nearest result plus the exact
 * rounding error (TwoSum / FMA residual) decides whether to step one ulp outward. */
static inline double add_rd(double a, double b)
{
    double s = a + b, bb = s - a, e = (a - (s - bb)) + (b - bb);
    return e < 0 ? nextafter(s, -INFINITY) : s;
}
static inline double add_ru(double a, double b)
{
    double s = a + b, bb = s - a, e = (a - (s - bb)) + (b - bb);
    return e > 0 ? nextafter(s, INFINITY) : s;
}
static inline double mul_rd(double a, double b)
{
    double p = a * b, e = fma(a, b, -p);
    return e < 0 ? nextafter(p, -INFINITY) : p;
}
static inline double mul_ru(double a, double b)
{
    double p = a * b, e = fma(a, b, -p);
    return e > 0 ? nextafter(p, INFINITY) : p;
}

/* Outward-rounded enclosure of the same recurrence: the lower chain rounds every product and sum
 * toward -inf, the upper chain toward +inf, and each chain feeds its own post-ReLU bounds forward.
 * Not in the reference (it rounds to nearest); required by the north star so that the reported
 * enclosure is sound.  By monotonicity lb_rd <= ws_lb and ub_ru >= ws_ub of fo_ibp. */
static void ibp_directed_one(const fo_net *m, const int32_t *box, double *scratch /* 4*maxw */,
                             double *lb_rd, double *ub_ru)
{
    double *lb = scratch, *ub = lb + m->maxw, *nlb = ub + m->maxw, *nub = nlb + m->maxw;
    int d = m->dims[0];
    for (int i = 0; i < d; ++i) { lb[i] = (double)box[2 * i]; ub[i] = (double)box[2 * i + 1]; }
    for (int l = 0; l < m->L; ++l) {
        int in = m->dims[l], out = m->dims[l + 1];
        const double *w = m->W + m->woff[l], *b = m->B + m->boff[l];
        for (int j = 0; j < out; ++j) {
            double mn = 0.0, mx = 0.0;
            for (int i = 0; i < in; ++i) {
                double wt = w[(int64_t)i * out + j];
                if (wt < 0) { mn = add_rd(mn, mul_rd(wt, ub[i])); mx = add_ru(mx, mul_ru(wt, lb[i])); }
                else        { mn = add_rd(mn, mul_rd(wt, lb[i])); mx = add_ru(mx, mul_ru(wt, ub[i])); }
            }
            double mnw = add_rd(mn, b[j]), mxw = add_ru(mx, b[j]);
            int64_t o = m->boff[l] + j;
            lb_rd[o] = mnw; ub_ru[o] = mxw;
            nlb[j] = mnw < 0 ? 0.0 : mnw; nub[j] = mxw < 0 ? 0.0 : mxw;
        }
        memcpy(lb, nlb, sizeof(double) * out); memcpy(ub, nub, sizeof(double) * out);
    }
}

void fo_ibp_directed(int L, const int32_t *dims, const double *W, const double *B, int64_t n,
                     const int32_t *boxes, double *lb_rd, double *ub_ru)
{
    fo_net m; if (fo_net_init(&m, L, dims, W, B)) return;
    int d = dims[0];
#pragma omp parallel
    {
        double *scratch = (double *)malloc(sizeof(double) * m.maxw * 4);
#pragma omp for schedule(static)
        for (int64_t p = 0; p < n; ++p)
            ibp_directed_one(&m, boxes + p * d * 2, scratch, lb_rd + p * m.N, ub_ru + p * m.N);
        free(scratch);
    }
}

/* ------------------------------------------------------------------------------------------------
 * Device-RNG restatement (throughput mode; the reference's MT19937 stream is injected from the host
 * in parity mode).  Philox4x32-10, counter = (sample, partition_lo, partition_hi, block),
 * key = (seed_lo, seed_hi); words are turned into bounded integers by fo_sample_row below.
 * Replaces utils/prune.py:205-222 (simluate_data: uniform integer in [lo, hi] per feature).
 * ---------------------------------------------------------------------------------------------- */
static inline void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                 uint32_t k0, uint32_t k1, uint32_t out[4])
{
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

/* Batched bounded integers: feature i takes the high half of word * range_i and hands the low half on as the
 * word of the next feature of its group.  G features share one Philox word, G chosen from the box's widest range so
 * that the product of the ranges drawn from one word stays <= 2^16: widest <= 16 -> 4, <= 256 -> 2, else 1.
 * Word q of a sample is output q % 4 of the Philox block with counter (sample, partition_lo, partition_hi, q / 4). */
static inline int fo_draw_group(int d, const int32_t *box)
{
    uint32_t widest = 0;
    for (int i = 0; i < d; ++i) {
        uint32_t range = (uint32_t)(box[2 * i + 1] - box[2 * i] + 1);
        if (range > widest) widest = range;
    }
    return widest <= 16u ? 4 : (widest <= 256u ? 2 : 1);
}

static inline void fo_sample_row(uint64_t seed, int64_t pidx, int s, int d, const int32_t *box, int32_t *x)
{
    const int G = fo_draw_group(d, box);
    uint32_t r[4] = {0, 0, 0, 0}, word = 0;
    for (int i = 0; i < d; ++i) {
        if (i % G == 0) {
            int q = i / G;
            if (q % 4 == 0)
                philox4x32_10((uint32_t)s, (uint32_t)pidx, (uint32_t)((uint64_t)pidx >> 32), (uint32_t)(q / 4),
                              (uint32_t)seed, (uint32_t)(seed >> 32), r);
            word = r[q % 4];
        }
        uint32_t range = (uint32_t)(box[2 * i + 1] - box[2 * i] + 1);
        uint64_t prod = (uint64_t)word * range;
        x[i] = box[2 * i] + (int32_t)(uint32_t)(prod >> 32);
        word = (uint32_t)prod;
    }
}

void fo_philox_samples(uint64_t seed, int64_t first_index, int64_t n, int S, int d,
                       const int32_t *boxes, int32_t *samples)
{
#pragma omp parallel for schedule(static)
    for (int64_t p = 0; p < n; ++p)
        for (int s = 0; s < S; ++s)
            fo_sample_row(seed, first_index + p, s, d, boxes + p * d * 2, samples + (p * S + s) * d);
}

/* ------------------------------------------------------------------------------------------------
 * a4 / a5: layer_net (utils/AC-1-Model-Functions.py:16-23, same body in all 53 files) and
 * candidate_dead_nodes (utils/prune.py:168-192).  x_{l+1} = relu(W_l^T x_l + b_l), last layer linear,
 * float64.  A neuron's count is the number of samples whose activation is != 0 (:179) -- for hidden
 * layers that is pre-activation > 0, for the linear output layer it is logit != 0.
 * ---------------------------------------------------------------------------------------------- */
static inline void fo_forward_counts(const fo_net *m, const double *x0, double *a, double *b2,
                                     uint32_t *counts, double *logit)
{
    const double *x = x0; double *y = a;
    for (int l = 0; l < m->L; ++l) {
        int in = m->dims[l], out = m->dims[l + 1];
        const double *w = m->W + m->woff[l], *b = m->B + m->boff[l];
        int last = (l == m->L - 1);
        for (int j = 0; j < out; ++j) {
            double acc = 0.0;
            for (int i = 0; i < in; ++i) acc = fma(w[(int64_t)i * out + j], x[i], acc);
            double pre = acc + b[j];
            double act = last ? pre : (pre > 0 ? pre : 0.0);
            if (counts && act != 0) counts[m->boff[l] + j] += 1;
            y[j] = act;
        }
        x = y; y = (y == a) ? b2 : a;
    }
    if (logit) *logit = x[0];
}

void fo_profile(int L, const int32_t *dims, const double *W, const double *B, int64_t n, int S,
                const int32_t *samples, uint32_t *counts)
{
    fo_net m; if (fo_net_init(&m, L, dims, W, B)) return;
    int d = dims[0];
#pragma omp parallel
    {
        double *buf = (double *)malloc(sizeof(double) * m.maxw * 3);
#pragma omp for schedule(static)
        for (int64_t p = 0; p < n; ++p) {
            uint32_t *c = counts + p * m.N;
            memset(c, 0, sizeof(uint32_t) * m.N);
            for (int s = 0; s < S; ++s) {
                const int32_t *xi = samples + (p * S + s) * d;
                for (int i = 0; i < d; ++i) buf[i] = (double)xi[i];
                fo_forward_counts(&m, buf, buf + m.maxw, buf + 2 * m.maxw, c, NULL);
            }
        }
        free(buf);
    }
}

/* a13: net (utils/AC-1-Model-Functions.py:25-32) -> pre-sigmoid logit, float64 inputs. */
void fo_forward_f64(int L, const int32_t *dims, const double *W, const double *B, int64_t mrows,
                    const double *X, double *logits)
{
    fo_net m; if (fo_net_init(&m, L, dims, W, B)) return;
    int d = dims[0];
#pragma omp parallel
    {
        double *buf = (double *)malloc(sizeof(double) * m.maxw * 3);
#pragma omp for schedule(static)
        for (int64_t r = 0; r < mrows; ++r) {
            memcpy(buf, X + r * d, sizeof(double) * d);
            fo_forward_counts(&m, buf, buf + m.maxw, buf + 2 * m.maxw, NULL, logits + r);
        }
        free(buf);
    }
}

/* float32 forward: what `net` computes when the driver feeds it a float32 point (C-check,
 * src/AC/Verify-AC.py:229-249) and what the pair screen's keras predict computes
 * (src/AC/metrics.py:323-325).  W32/B32 are the original float32 weights. */
static inline float fo_logit_f32(int L, const int32_t *dims, const float *W32, const float *B32,
                                 const float *x0, float *a, float *b2)
{
    const float *x = x0; float *y = a; int64_t wo = 0, bo = 0;
    for (int l = 0; l < L; ++l) {
        int in = dims[l], out = dims[l + 1];
        const float *w = W32 + wo, *b = B32 + bo;
        int last = (l == L - 1);
        for (int j = 0; j < out; ++j) {
            float acc = 0.0f;
            for (int i = 0; i < in; ++i) acc = fmaf(w[(int64_t)i * out + j], x[i], acc);
            float pre = acc + b[j];
            y[j] = last ? pre : (pre > 0 ? pre : 0.0f);
        }
        wo += (int64_t)in * out; bo += out;
        x = y; y = (y == a) ? b2 : a;
    }
    return x[0];
}

static int fo_maxw(int L, const int32_t *dims)
{
    int mw = dims[0];
    for (int l = 1; l <= L; ++l) if (dims[l] > mw) mw = dims[l];
    return mw;
}

void fo_forward_f32(int L, const int32_t *dims, const float *W32, const float *B32, int64_t mrows,
                    const float *X, float *logits)
{
    int d = dims[0], mw = fo_maxw(L, dims);
#pragma omp parallel
    {
        float *buf = (float *)malloc(sizeof(float) * mw * 3);
#pragma omp for schedule(static)
        for (int64_t r = 0; r < mrows; ++r) {
            memcpy(buf, X + r * d, sizeof(float) * d);
            logits[r] = fo_logit_f32(L, dims, W32, B32, buf, buf + mw, buf + 2 * mw);
        }
        free(buf);
    }
}

/* ------------------------------------------------------------------------------------------------
 * a9: singular_verification*, utils/prune.py:276-644, as the closed form of the query the reference
 * really poses to Z3.  For a still-candidate neuron j of layer l (layers 0..L-2) it asks whether
 * {input constraints} and y_j > 0 is satisfiable, y = W_l^T x + b_l, and marks the neuron dead on unsat.
 *   layer 0 (:297-300): x are Int variables inside the partition box (input_domain_constraint :253-263).
 *     An affine function attains its supremum over an integer box at a corner, so the query is sat iff
 *     the exact value of ws_ub[0][j] is > 0.
 *   layer l >= 1 (:302-340): x are in_l Real variables and the constraints come from
 *     intermediate_domain_constraint(x, pl_lb, pl_ub, layer_index) (:266-273) -- called with the layer's
 *     OWN index, so x_k is boxed by pl_lb[l][k] <= x_k <= pl_ub[l][k] for k < out_l (the post-activation
 *     bounds of layer l itself, index for index) and the inputs k >= out_l stay unbounded.  (For
 *     out_l > in_l the reference raises IndexError at x[neuron_index]; no reference model has such a
 *     layer.  Here the constraints stop at min(in_l, out_l).)  The query is therefore sat iff some
 *     unbounded input has a non-zero weight, or
 *         q_j = sum_{k < min(in_l, out_l)} max(w_kj * pl_lb[l][k], w_kj * pl_ub[l][k]) + b_j  >  0
 *     in exact arithmetic.  That is NOT the interval upper bound ws_ub[l][j]: on layers with in_l == out_l
 *     the reference declares neurons dead whose ws_ub is far above zero (AC-10, src grid partition 13529,
 *     layer 1 neuron 1: ws_ub = 14.47, q = -10.13, s_dead = 1).  Parity means reproducing exactly that.
 * q is evaluated like neuron_bounds evaluates its sums: k ascending, product and sum rounded separately,
 * bias last (Z3 itself is exact, on the decimal repr of every float; the difference is covered by tol).
 * q[] and unb[] are [N]; entries of layer 0 are (ws_ub, 0), of the last layer (0, 0).
 * ---------------------------------------------------------------------------------------------- */
static void singular_one(const fo_net *m, const double *ws_ub, const double *pl_lb, const double *pl_ub,
                         double *q, uint8_t *unb)
{
    for (int l = 0; l < m->L; ++l) {
        int in = m->dims[l], out = m->dims[l + 1];
        const double *w = m->W + m->woff[l], *b = m->B + m->boff[l];
        const double *xl = pl_lb + m->boff[l], *xu = pl_ub + m->boff[l];
        int mm = in < out ? in : out;
        for (int j = 0; j < out; ++j) {
            int64_t o = m->boff[l] + j;
            unb[o] = 0;
            if (l == m->L - 1) { q[o] = 0.0; continue; }
            if (l == 0) { q[o] = ws_ub[o]; continue; }
            double acc = 0.0;
            for (int k = 0; k < mm; ++k) {
                double wt = w[(int64_t)k * out + j];
                acc += (wt < 0) ? wt * xl[k] : wt * xu[k];
            }
            q[o] = acc + b[j];
            for (int k = out; k < in; ++k) if (w[(int64_t)k * out + j] != 0.0) unb[o] = 1;
        }
    }
}

/* ------------------------------------------------------------------------------------------------
 * a7 / a8 / a10: dead_node_from_bound (utils/prune.py:226-251), the keep-one fix
 * (:689-691, :698-700, :703-705) and merge_dead_nodes (:941-948), for a batch of partitions.
 *   cand   = count == 0                                   (:183-187, every layer incl. output)
 *   b_dead = cand & (ws_ub <= 0) on layers 0..L-2, = cand on the last layer (:235 break), keep-one
 *   b_cand = cand & !(ws_ub <= 0) on layers 0..L-2, = cand on the last layer
 *   s_dead = b_cand & !unb & (q <= 0) on layers 0..L-2 (singular_one above; on layer 0 q = ws_ub > 0, so 0),
 *            = b_cand on the last layer (:287-288 break leaves the copy), keep-one
 *   s_cand = b_cand minus the neurons just found dead (before the keep-one fix, as in :349-350)
 *   sliver = b_cand & !unb & |q| <= tol[j]: the float sign of q is within rounding error of the exact one;
 *            the host runs the reference's own Z3 query for these and overrides s_dead / s_cand / dead
 *   dead   = b_dead | s_dead, keep-one
 * All masks are bytes [n][N].
 * ---------------------------------------------------------------------------------------------- */
static void keep_one(uint8_t *mask, const fo_net *m)
{
    for (int l = 0; l < m->L; ++l) {
        int out = m->dims[l + 1]; uint8_t *q = mask + m->boff[l]; int has0 = 0;
        for (int j = 0; j < out; ++j) if (!q[j]) { has0 = 1; break; }
        if (!has0) q[0] = 0;
    }
}

static void masks_one(const fo_net *m, const uint32_t *counts, const double *ws_ub, const double *q,
                      const uint8_t *unb, const double *tol, uint8_t *cand, uint8_t *b_dead, uint8_t *b_cand,
                      uint8_t *s_dead, uint8_t *s_cand, uint8_t *dead, uint8_t *sliver)
{
    for (int l = 0; l < m->L; ++l)
        for (int j = 0; j < m->dims[l + 1]; ++j) {
            int64_t k = m->boff[l] + j;
            uint8_t c = counts ? counts[k] == 0 : 1;
            cand[k] = c; sliver[k] = 0;
            if (l == m->L - 1) { b_dead[k] = c; b_cand[k] = c; s_dead[k] = c; s_cand[k] = c; }
            else {
                uint8_t le0 = ws_ub[k] <= 0;
                b_dead[k] = c & le0; b_cand[k] = c & !le0;
                s_dead[k] = b_cand[k] && !unb[k] && q[k] <= 0;
                s_cand[k] = b_cand[k] & !s_dead[k];
                if (tol && b_cand[k] && !unb[k] && fabs(q[k]) <= tol[k]) sliver[k] = 1;
            }
        }
    keep_one(b_dead, m);
    keep_one(s_dead, m);
    for (int k = 0; k < m->N; ++k) dead[k] = b_dead[k] | s_dead[k];
    keep_one(dead, m);
}

void fo_masks(int L, const int32_t *dims, const double *W, const double *B, int64_t n, const uint32_t *counts,
              const double *ws_ub, const double *pl_lb, const double *pl_ub, const double *tol, uint8_t *cand,
              uint8_t *b_dead, uint8_t *b_cand, uint8_t *s_dead, uint8_t *s_cand, uint8_t *dead, uint8_t *sliver,
              double *q_out)
{
    fo_net m; if (fo_net_init(&m, L, dims, W, B)) return;
#pragma omp parallel
    {
        double *q = (double *)malloc(sizeof(double) * m.N);
        uint8_t *unb = (uint8_t *)malloc(m.N);
#pragma omp for schedule(static)
        for (int64_t p = 0; p < n; ++p) {
            int64_t o = p * m.N;
            singular_one(&m, ws_ub + o, pl_lb + o, pl_ub + o, q, unb);
            masks_one(&m, counts ? counts + o : NULL, ws_ub + o, q, unb, tol, cand + o, b_dead + o, b_cand + o,
                      s_dead + o, s_cand + o, dead + o, sliver + o);
            if (q_out) memcpy(q_out + o, q, sizeof(double) * m.N);
        }
        free(q); free(unb);
    }
}

/* compression_ratio, utils/prune.py:194-203: 1 - alive/total over all layers incl. the output. */
double fo_compression_ratio(int N, const uint8_t *mask)
{
    int alive = 0;
    for (int k = 0; k < N; ++k) alive += !mask[k];
    return 1.0 - ((double)alive / (double)N);
}

/* ------------------------------------------------------------------------------------------------
 * a12: heuristic_prune, utils/prune.py:862-939, with numpy's float64 mean / median / percentile
 * restated (numpy 2.x: add.reduce = pairwise sum, 8 accumulators per block of <= 128; percentile = linear
 * method, virtual index (n-1)*q, lerp with the t >= 0.5 branch); verified bit-for-bit against numpy in
 * tests/test_oracle_golden.py::test_numpy_statistics_restatement.
 * ---------------------------------------------------------------------------------------------- */
static double np_pairwise(const double *a, int n)
{
    if (n < 8) { double r = 0.0; for (int i = 0; i < n; ++i) r += a[i]; return r; }
    if (n <= 128) {
        double r[8]; int i;
        for (i = 0; i < 8; ++i) r[i] = a[i];
        for (i = 8; i < n - (n % 8); i += 8) for (int k = 0; k < 8; ++k) r[k] += a[i + k];
        double res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
        for (; i < n; ++i) res += a[i];
        return res;
    }
    int n2 = n / 2; n2 -= n2 % 8;
    return np_pairwise(a, n2) + np_pairwise(a + n2, n - n2);
}
double fo_np_sum(const double *a, int n) { return np_pairwise(a, n); }
double fo_np_mean(const double *a, int n) { return fo_np_sum(a, n) / (double)n; }

static int cmp_double(const void *x, const void *y)
{
    double a = *(const double *)x, b = *(const double *)y;
    return (a > b) - (a < b);
}
/* `sorted` must be ascending. */
double fo_np_median_sorted(const double *sorted, int n)
{
    if (n & 1) return sorted[n / 2];
    double pair[2] = { sorted[n / 2 - 1], sorted[n / 2] };
    return fo_np_mean(pair, 2);
}
double fo_np_percentile_sorted(const double *sorted, int n, double pct)
{
    double q = pct / 100.0;
    double v = (double)(n - 1) * q;
    int64_t prev, next; double gamma;
    if (v >= (double)(n - 1)) { prev = next = n - 1; gamma = v - (double)(n - 1); }
    else if (v < 0)           { prev = next = 0; gamma = v - 0.0; }
    else { prev = (int64_t)floor(v); next = prev + 1; gamma = v - (double)prev; }
    double a = sorted[prev], b = sorted[next], diff = b - a;
    double r = a + diff * gamma;
    if (gamma >= 0.5) r = b - diff * (1.0 - gamma);
    return r;
}

/* One partition.  cand / s_cand / deads are byte masks [N]; ws_lb / ws_ub are [N]. */
void fo_heuristic_prune(int L, const int32_t *dims, const double *ws_lb, const double *ws_ub,
                        const uint8_t *cand, const uint8_t *s_cand, const uint8_t *deads,
                        double pct, uint8_t *new_deads, uint8_t *merged)
{
    fo_net m; if (fo_net_init(&m, L, dims, NULL, NULL)) return;
    memset(new_deads, 0, m.N);
    double *cv = (double *)malloc(sizeof(double) * m.maxw * 2), *nv = cv + m.maxw;
    for (int l = 0; l + 1 < L; ++l) {
        int out = dims[l + 1], nc = 0, nn = 0; int64_t o = m.boff[l];
        for (int j = 0; j < out; ++j) { if (cand[o + j]) cv[nc++] = ws_ub[o + j]; else nv[nn++] = ws_ub[o + j]; }
        if (nn == 0) { for (int j = 0; j < out; ++j) new_deads[o + j] = 1; continue; }
        if (nc == 0) continue;
        double cmean = fo_np_mean(cv, nc), nmean = fo_np_mean(nv, nn);
        qsort(cv, nc, sizeof(double), cmp_double); qsort(nv, nn, sizeof(double), cmp_double);
        double cmed = fo_np_median_sorted(cv, nc), nmed = fo_np_median_sorted(nv, nn);
        double n_lo = fo_np_percentile_sorted(nv, nn, pct), n_hi = fo_np_percentile_sorted(nv, nn, 100.0 - pct);
        if (nmean > 2 * cmean && nmed > 2 * cmed)
            for (int j = 0; j < out; ++j)
                if (s_cand[o + j] && ws_ub[o + j] < n_lo && ws_ub[o + j] < 0.1 * n_hi &&
                    ws_ub[o + j] < fabs(ws_lb[o + j]))
                    new_deads[o + j] = 1;
    }
    free(cv);
    keep_one(new_deads, &m);
    for (int k = 0; k < m.N; ++k) merged[k] = deads[k] | new_deads[k];
    keep_one(merged, &m);
}

/* ------------------------------------------------------------------------------------------------
 * a15: the prediction side of CausalDiscriminationDetector.causal_discrimination
 * (src/AC/metrics.py:130-159).  The sampler's RNG stream does not depend on predictions, so the host
 * draws every base row first; this evaluates, per row, the original class and the index of the first
 * alternative protected assignment (in itertools.product order, :219-227, skipping the row's own
 * assignment, :147-148) whose class differs (:155-159), or -1.  Class = float32 logit > 0
 * (sigmoid(logit) > 0.5, :325).
 * ---------------------------------------------------------------------------------------------- */
void fo_pair_screen(int L, const int32_t *dims, const float *W32, const float *B32, int64_t mrows,
                    const float *X, int n_pa, const int32_t *pa_cols, int k_alts, const float *alts,
                    uint8_t *orig_pred, int32_t *first_flip)
{
    int d = dims[0], mw = fo_maxw(L, dims);
#pragma omp parallel
    {
        float *buf = (float *)malloc(sizeof(float) * (mw * 3 + d));
        float *x = buf + mw * 3;
#pragma omp for schedule(static)
        for (int64_t r = 0; r < mrows; ++r) {
            memcpy(x, X + r * d, sizeof(float) * d);
            memcpy(buf, x, sizeof(float) * d);
            int p0 = fo_logit_f32(L, dims, W32, B32, buf, buf + mw, buf + 2 * mw) > 0;
            orig_pred[r] = (uint8_t)p0; first_flip[r] = -1;
            for (int a = 0; a < k_alts; ++a) {
                int same = 1;
                for (int q = 0; q < n_pa; ++q) if (alts[a * n_pa + q] != X[r * d + pa_cols[q]]) same = 0;
                if (same) continue;
                memcpy(buf, x, sizeof(float) * d);
                for (int q = 0; q < n_pa; ++q) buf[pa_cols[q]] = alts[a * n_pa + q];
                int p1 = fo_logit_f32(L, dims, W32, B32, buf, buf + mw, buf + 2 * mw) > 0;
                if (p1 != p0) { first_flip[r] = a; break; }
            }
        }
        free(buf);
    }
}

/* ------------------------------------------------------------------------------------------------
 * Per-partition sampled counterexample screen (north star; no counterpart function in the reference,
 * it restates the property the drivers hand to Z3, src/AC/Verify-AC.py:137-156: two points of the
 * partition that differ only in the protected attribute and whose logits have strictly opposite
 * signs).  For each sample the protected attribute is swept over the partition's range [lo, hi] of
 * that column; a sample is a counterexample if some value gives logit > 0 and another logit < 0
 * (float32 forward, as keras predicts).  Reported: the lowest such sample index, the smallest value
 * with a positive logit and the smallest with a negative one; -1 when none.
 * ---------------------------------------------------------------------------------------------- */
void fo_partition_screen(int L, const int32_t *dims, const float *W32, const float *B32, int64_t n,
                         int S, const int32_t *samples, const int32_t *boxes, int pa_col,
                         int32_t *ce_sample, int32_t *pa_pos, int32_t *pa_neg)
{
    int d = dims[0], mw = fo_maxw(L, dims);
#pragma omp parallel
    {
        float *buf = (float *)malloc(sizeof(float) * (mw * 3 + d));
        float *x = buf + mw * 3;
#pragma omp for schedule(static)
        for (int64_t p = 0; p < n; ++p) {
            int lo = boxes[(p * d + pa_col) * 2], hi = boxes[(p * d + pa_col) * 2 + 1];
            ce_sample[p] = -1; pa_pos[p] = -1; pa_neg[p] = -1;
            for (int s = 0; s < S && ce_sample[p] < 0; ++s) {
                const int32_t *xi = samples + (p * S + s) * d;
                for (int i = 0; i < d; ++i) x[i] = (float)xi[i];
                int vp = -1, vn = -1, havep = 0, haven = 0;
                for (int v = lo; v <= hi; ++v) {
                    memcpy(buf, x, sizeof(float) * d);
                    buf[pa_col] = (float)v;
                    float y = fo_logit_f32(L, dims, W32, B32, buf, buf + mw, buf + 2 * mw);
                    if (y > 0 && !havep) { havep = 1; vp = v; }
                    if (y < 0 && !haven) { haven = 1; vn = v; }
                }
                if (havep && haven) { ce_sample[p] = s; pa_pos[p] = vp; pa_neg[p] = vn; }
            }
        }
        free(buf);
    }
}

/* ------------------------------------------------------------------------------------------------
 * The same screen with the RELAXED query of relaxed/AC/Verify-AC.py:48-51, :128-132: the two points may also differ
 * in column ra_col by at most thr (in_const_diff_adult, utils/verif_utils.py:859-867).  Only the first point has to
 * lie in the partition (in_const_domain_adult bounds x' on the protected columns alone, :846-854), so the second
 * point's ra value ranges over a - thr .. a + thr around the sample's value a, unclipped.  Per sample: the strict
 * sweep first (smallest protected value with a positive logit, smallest with a negative one); if it shows one class
 * only, the first second point -- difference d ascending from -thr (d != 0), protected value ascending -- of the
 * other class that has a partner with a different protected value in the strict sweep.  ra_col < 0: strict only.
 * ra_dpos / ra_dneg: the difference applied to the positive / negative point (0 for the one inside the box).
 * ---------------------------------------------------------------------------------------------- */
void fo_partition_screen_ex(int L, const int32_t *dims, const float *W32, const float *B32, int64_t n,
                            int S, const int32_t *samples, const int32_t *boxes, int pa_col, int ra_col, int thr,
                            int32_t *ce_sample, int32_t *pa_pos, int32_t *pa_neg, int32_t *ra_dpos, int32_t *ra_dneg)
{
    int d = dims[0], mw = fo_maxw(L, dims);
#pragma omp parallel
    {
        float *buf = (float *)malloc(sizeof(float) * (mw * 3 + d));
        float *x = buf + mw * 3;
#pragma omp for schedule(static)
        for (int64_t p = 0; p < n; ++p) {
            int lo = boxes[(p * d + pa_col) * 2], hi = boxes[(p * d + pa_col) * 2 + 1];
            ce_sample[p] = -1; pa_pos[p] = -1; pa_neg[p] = -1; ra_dpos[p] = 0; ra_dneg[p] = 0;
            for (int s = 0; s < S && ce_sample[p] < 0; ++s) {
                const int32_t *xi = samples + (p * S + s) * d;
                for (int i = 0; i < d; ++i) x[i] = (float)xi[i];
                int vp = -1, vn = -1, vp2 = -1, vn2 = -1;
                for (int v = lo; v <= hi; ++v) {
                    memcpy(buf, x, sizeof(float) * d);
                    buf[pa_col] = (float)v;
                    float y = fo_logit_f32(L, dims, W32, B32, buf, buf + mw, buf + 2 * mw);
                    if (y > 0) { if (vp < 0) vp = v; else if (vp2 < 0) vp2 = v; }
                    if (y < 0) { if (vn < 0) vn = v; else if (vn2 < 0) vn2 = v; }
                }
                if (vp >= 0 && vn >= 0) { ce_sample[p] = s; pa_pos[p] = vp; pa_neg[p] = vn; continue; }
                if (ra_col < 0 || (vp < 0) == (vn < 0)) continue;
                for (int dd = -thr; dd <= thr && ce_sample[p] < 0; ++dd) {
                    if (dd == 0) continue;
                    for (int v = lo; v <= hi && ce_sample[p] < 0; ++v) {
                        memcpy(buf, x, sizeof(float) * d);
                        buf[pa_col] = (float)v; buf[ra_col] = (float)(xi[ra_col] + dd);
                        float y = fo_logit_f32(L, dims, W32, B32, buf, buf + mw, buf + 2 * mw);
                        if (vp >= 0 && y < 0) {
                            int partner = (vp != v) ? vp : vp2;
                            if (partner >= 0) { ce_sample[p] = s; pa_pos[p] = partner; pa_neg[p] = v; ra_dneg[p] = dd; }
                        } else if (vn >= 0 && y > 0) {
                            int partner = (vn != v) ? vn : vn2;
                            if (partner >= 0) { ce_sample[p] = s; pa_pos[p] = v; pa_neg[p] = partner; ra_dpos[p] = dd; }
                        }
                    }
                }
            }
        }
        free(buf);
    }
}

/* ------------------------------------------------------------------------------------------------
 * Diagnostic of the summation-order question (see fb_profile_nearzero in include/fairify_b200.h): per (sample, neuron)
 * the ratio |pre| / (sum |w x| + |b|); count of ratios <= 2^-40 and the smallest ratio.
 * ---------------------------------------------------------------------------------------------- */
void fo_nearzero(int L, const int32_t *dims, const double *W, const double *B, int64_t n, int S, const int32_t *samples,
                 uint64_t *count, double *min_ratio)
{
    fo_net m; if (fo_net_init(&m, L, dims, W, B)) return;
    int d = dims[0];
    uint64_t cnt = 0; double best = 1.0;
    double *a = (double *)malloc(sizeof(double) * m.maxw * 2), *y = a + m.maxw;
    for (int64_t u = 0; u < n * S; ++u) {
        for (int i = 0; i < d; ++i) a[i] = (double)samples[u * d + i];
        double *in = a, *out = y;
        for (int l = 0; l < L; ++l) {
            int IN = dims[l], OUT = dims[l + 1];
            const double *w = W + m.woff[l], *b = B + m.boff[l];
            for (int j = 0; j < OUT; ++j) {
                double acc = 0.0, mag = 0.0;
                for (int i = 0; i < IN; ++i) { acc = fma(w[(int64_t)i * OUT + j], in[i], acc); mag = fma(fabs(w[(int64_t)i * OUT + j]), fabs(in[i]), mag); }
                double pre = acc + b[j]; mag = mag + fabs(b[j]);
                if (mag > 0.0) { double r = fabs(pre) / mag; if (r <= 0x1p-40) ++cnt; if (r < best) best = r; }
                out[j] = (l == L - 1) ? pre : (pre > 0.0 ? pre : 0.0);
            }
            double *t = in; in = out; out = t;
        }
    }
    free(a);
    *count = cnt; *min_ratio = best;
}

/* ------------------------------------------------------------------------------------------------
 * Whole pre-solver for a batch of boxes (a3 Philox + a5 + a6 + a7 + a8 + a10), used as the timed CPU
 * baseline and as the end-to-end checker of the fused device path.  Outputs: counts [n][N] and the
 * byte masks of fo_masks, plus the verdict byte: bit0 = output logit provably > 0 or < 0 over the
 * whole box by the outward-rounded enclosure (partition certified fair without Z3), bit2 = some
 * neuron is a rounding sliver.
 * ---------------------------------------------------------------------------------------------- */
void fo_prune_batch(int L, const int32_t *dims, const double *W, const double *B, int64_t n,
                    const int32_t *boxes, uint64_t seed, int64_t first_index, int S, const double *tol,
                    uint32_t *counts, uint8_t *cand, uint8_t *b_dead, uint8_t *s_dead, uint8_t *dead,
                    uint8_t *sliver, uint8_t *verdict)
{
    fo_net m; if (fo_net_init(&m, L, dims, W, B)) return;
    int d = dims[0];
#pragma omp parallel
    {
        double *buf = (double *)malloc(sizeof(double) * (m.maxw * 4 + 7 * m.N));
        double *wl = buf + m.maxw * 4, *wu = wl + m.N, *pl = wu + m.N, *pu = pl + m.N, *rd = pu + m.N, *ru = rd + m.N;
        double *q = ru + m.N;
        int32_t *xi = (int32_t *)malloc(sizeof(int32_t) * d);
        uint8_t *bc = (uint8_t *)malloc(3 * m.N), *sc = bc + m.N, *unb = sc + m.N;
#pragma omp for schedule(static)
        for (int64_t p = 0; p < n; ++p) {
            const int32_t *box = boxes + p * d * 2;
            uint32_t *c = counts + p * m.N;
            memset(c, 0, sizeof(uint32_t) * m.N);
            for (int s = 0; s < S; ++s) {
                fo_sample_row(seed, first_index + p, s, d, box, xi);
                for (int i = 0; i < d; ++i) buf[i] = (double)xi[i];
                fo_forward_counts(&m, buf, buf + m.maxw, buf + 2 * m.maxw, c, NULL);
            }
            ibp_one(&m, box, buf, wl, wu, pl, pu);
            ibp_directed_one(&m, box, buf, rd, ru);
            singular_one(&m, wu, pl, pu, q, unb);
            masks_one(&m, c, wu, q, unb, tol, cand + p * m.N, b_dead + p * m.N, bc, s_dead + p * m.N, sc,
                      dead + p * m.N, sliver + p * m.N);
            uint8_t v = 0;
            if (rd[m.N - 1] > 0 || ru[m.N - 1] < 0) v |= 1;
            for (int k = 0; k < m.N; ++k) if (sliver[p * m.N + k]) v |= 4;
            verdict[p] = v;
        }
        free(buf); free(xi); free(bc);
    }
}
