/*
 * fairify_b200.h -- C ABI of libfairify_b200.so: Fairify's data-parallel pre-solver stage on B200 (sm_100a).
 *
 * Plain pointers and sizes only.  Unless a function name ends in _host, every data pointer is a DEVICE
 * pointer on the device that was current when the model was created, and `stream` is a cudaStream_t
 * passed as void* (NULL = the legacy default stream).  Functions return 0 on success, a non-zero code on
 * failure; fb_last_error() then returns a thread-local description.  There is no CPU fallback: without
 * a usable CUDA device every compute entry point fails.
 *
 * Each entry point names the reference interface it replaces (paths relative to the Fairify checkout).
 *
 * Layouts
 *   weights   float32, layer-major, each layer [in][out] row-major -- exactly what
 *             `model.layers[i].get_weights()` returns in the drivers (src/AC/Verify-AC.py:97-99)
 *   boxes     int32 [n][d][2] = (lo, hi) per feature in DataFrame COLUMN order
 *             (the order utils/prune.py:108-114 looks the range_dict up in)
 *   samples   int32 [n][S][d], or NULL to draw them on the device: Philox4x32-10, counter = (sample,
 *             partition index, block), key = seed; each 32-bit word yields 4, 2 or 1 bounded integers (multiply-
 *             shift, the low half of the product handed on) depending on whether the box's widest range is
 *             <= 16, <= 256 or larger, so every value is uniform to within 2^-16 relative.  A feature's range
 *             hi - lo + 1 must fit in 32 bits (hi - lo < 2^32 - 1)
 *   masks     uint32 [n][NW], NW = ceil(N/32); bit k = neuron k, neurons numbered layer by layer,
 *             output neuron included (N = sum of layer widths)
 *   per-neuron arrays  [n][N] in the same numbering
 */
#ifndef FAIRIFY_B200_H
#define FAIRIFY_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FB_OK 0
#define FB_ERR_INVALID 1      /* bad argument */
#define FB_ERR_CUDA 2         /* CUDA runtime error (message has the cudaError string) */
#define FB_ERR_UNSUPPORTED 3  /* shape outside the supported envelope */
#define FB_ERR_NOMEM 4

#define FB_VERDICT_CERTIFIED 1u /* logit provably > 0 or < 0 on the whole box: no counterexample exists */
#define FB_VERDICT_CE 2u        /* the sampled screen found a counterexample pair */
#define FB_VERDICT_SLIVER 4u    /* a bound is within rounding error of 0: host must run the Z3 neuron check */
#define FB_VERDICT_SKIPPED 8u   /* the partition screen skipped this box: its protected range does not fit the limits below */

typedef struct fb_model fb_model;
typedef struct fb_grid fb_grid;

/* Outputs of the prune stage; any pointer may be NULL (that output is skipped). */
typedef struct fb_prune_out {
    uint32_t *cand;     /* candidates: never active on the samples      (candidate_dead_nodes, utils/prune.py:183-187) */
    uint32_t *b_dead;   /* b_dead_node_mask after the keep-one fix       (dead_node_from_bound :226-251, :689-691)     */
    uint32_t *dead;     /* dead_nodes = merge(b_dead, s_dead) + keep-one (:702-705)                                    */
    uint32_t *sliver;   /* neurons whose s_dead the float arithmetic cannot decide: hand them to the Z3 query (:276-356) */
    uint8_t *verdict;   /* [n] FB_VERDICT_* bits                                                                       */
    uint16_t *counts;   /* [n][N] non-zero activation counts; pos_prob = counts / S (:188-191)                         */
    double *ws_lb, *ws_ub, *pl_lb, *pl_ub; /* [n][N] neuron_bounds, round-to-nearest, reference-exact (:105-164)      */
    double *lb_rd, *ub_ru;                 /* [n][N] outward-rounded enclosure of the same recurrence                  */
    uint32_t *s_dead;   /* s_dead_node_mask of singular_verification* (:276-644) after the keep-one fix (:698-700), from
                         * the closed form of the query the reference gives Z3 (see fb_ibp)                            */
    uint32_t *s_cand;   /* s_candidates: candidates neither the bound nor the query settled (:349-350)                 */
    uint32_t *record;   /* [n][2*NW+2] packed per-partition record for the multi-GPU gather: verdict | dead[NW] |
                         * cand[NW] | first counterexample sample (written as -1; fb_partition_screen_ex fills it)     */
} fb_prune_out;

/* -- library ---------------------------------------------------------------------------------- */
int fb_version(void);
const char *fb_last_error(void);
/* number of SMs of the current device (grid sizing; 148 on B200) */
int fb_num_sms(int *out);

/* -- model: replaces the (w, b) lists every hot-path function receives -------------------------- */
/* L weight layers, dims[L+1] = {d, h1, ..., 1}.  Weights are converted to float64 once (exact). */
int fb_model_create(int L, const int32_t *dims, const float *w_host, const float *b_host, fb_model **out);
void fb_model_destroy(fb_model *m);
/* N = neurons, P = weights, specialised = 1 when a compile-time-unrolled kernel set serves this shape */
int fb_model_info(const fb_model *m, int *L, int *N, int *P, int *specialised);
/* the device the model's weights live on; every call that takes the model fails unless that device is current */
int fb_model_device(const fb_model *m, int *device);
/* fb_prune, fb_prune_host and fb_partition_screen draw their scratch and staging buffers from a memory pool private to
 * this library (one per device) and keep freed blocks cached there for the next call; this returns the cached memory
 * to the driver (call it between phases, or before another component needs the whole GPU). */
int fb_release_cache(void);

/* -- partition grid: utils/input_partition.py:17-46 `partition`, :48-76 `partitioned_ranges` ---- */
/* lo/hi in column order, T = PARTITION_THRESHOLD, key_order[k] = column index of the k-th range_dict key */
int fb_grid_create(int d, const int64_t *lo, const int64_t *hi, int64_t T, const int32_t *key_order, fb_grid **out);
void fb_grid_destroy(fb_grid *g);
int64_t fb_grid_size(const fb_grid *g);
/* boxes [count][d][2] for canonical (itertools.product) indices first .. first+count-1 */
int fb_grid_boxes(const fb_grid *g, int64_t first, int64_t count, int32_t *boxes, void *stream);

/* point -> partition: canonical index of the box containing each row of X (float64 [m][d], column order), -1 if none;
 * replaces the linear scans of find_partition_result_for_point (src/AC/Verify-AC-experiment-new2.py:587-611) and
 * p_list_density (utils/input_partition.py:184-218).  index int64 [m] and/or density uint32 [grid size] (rows per
 * partition, ADDED to what is there: zero it first); either may be NULL. */
int fb_grid_locate(const fb_grid *g, int64_t m, const double *X, int64_t *index, uint32_t *density, void *stream);

/* -- activation profile: candidate_dead_nodes, utils/prune.py:168-192 (+ simluate_data :205-222) -- */
/* counts uint32 [n][N].  first_index = global index of boxes[0] (Philox counter; keeps shards identical) */
int fb_profile(const fb_model *m, int64_t n, const int32_t *boxes, const int32_t *samples, int S, uint64_t seed,
               int64_t first_index, uint32_t *counts, void *stream);

/* -- interval bounds + masks: neuron_bounds :105-164, dead_node_from_bound :226-251, singular_verification*
 *    :276-644, keep-one fixes, merge_dead_nodes :941-948.  counts may be NULL (bounds-only run: every neuron is a
 *    candidate).
 *    Singular verification runs on the device in closed form.  For a remaining candidate j of hidden layer l >= 1 the
 *    reference's Z3 query boxes input k < out_l by the layer's OWN post-activation bounds pl_lb[l][k], pl_ub[l][k]
 *    (intermediate_domain_constraint :266-273 is called with layer_index, :302-345) and leaves inputs k >= out_l free, so
 *    it is unsat -- s_dead = 1 -- iff no free input has a non-zero weight and
 *        q_j = sum_{k < min(in_l, out_l)} max(w_kj pl_lb[l][k], w_kj pl_ub[l][k]) + b_j <= 0.
 *    On layer 0 (integer box) the quantity is ws_ub itself.  tol [N] (may be NULL) is the caller's bound on the rounding
 *    error of that quantity per neuron; neurons with |q_j| <= tol_j are reported in `sliver` so the host can put the
 *    exact question to Z3.  (For out_l > in_l at l >= 1 the reference raises IndexError; here the sum stops at in_l.) */
int fb_ibp(const fb_model *m, int64_t n, const int32_t *boxes, const uint32_t *counts, const double *tol,
           const fb_prune_out *out, void *stream);

/* -- the whole pre-solver stage for a batch of boxes: sound_prune*, utils/prune.py:671-859; the Z3 neuron
 *    queries are answered in closed form (fb_ibp), the undecidable ones reported through `sliver`.
 *    Allocates [n][N] uint32 of scratch per call. */
int fb_prune(const fb_model *m, int64_t n, const int32_t *boxes, const int32_t *samples, int S, uint64_t seed,
             int64_t first_index, const double *tol, const fb_prune_out *out, void *stream);

/* Same with HOST buffers in and out (boxes, samples, tol and every non-NULL output are host pointers).
 * The batch is processed in slices of 10^6 partitions through a three-stage pipeline -- upload, kernels (on
 * `stream`), download -- on two internal copy streams, so with pinned host memory the copies hide behind the
 * kernels; results do not depend on the slicing.  Returns after everything has finished. */
int fb_prune_host(const fb_model *m, int64_t n, const int32_t *boxes, const int32_t *samples, int S, uint64_t seed,
                  int64_t first_index, const double *tol, const fb_prune_out *out, void *stream);

/* -- fb_prune for callers that must not allocate.  (No counterpart in the reference, which never manages
 *    device memory; it exists for embedders of this library.)  Same arguments, results and error
 *    convention as fb_prune, but the [n][N] uint32 count scratch is a fixed 16 MiB array inside the library
 *    image instead of a per-call allocation, so the call issues no cudaMalloc / cudaFree of any kind.
 *    Meant for hosts that forbid device allocation after start-up (allocator locks stall other streams),
 *    that have already carved up all of HBM, or that account every allocation themselves.
 *    Limits: n * N <= 4 194 304 (FB_ERR_UNSUPPORTED beyond that; e.g. 167 772 partitions for AC-1).
 *    The first call on each device resolves the array's address; make it at start-up, before any stream
 *    capture.  All calls on a device share the one array: calls that overlap on the GPU race. */
int fb_prune_noalloc(const fb_model *m, int64_t n, const int32_t *boxes, const int32_t *samples, int S, uint64_t seed,
                     int64_t first_index, const double *tol, const fb_prune_out *out, void *stream);

/* -- logits: `net`, utils/<MODEL>-Model-Functions.py (e.g. AC-1 :25-32); get_y_pred utils/verif_utils.py:1040-1047.
 *    dead_bits [NW] or NULL: neurons forced to 0 (== the pruned network of prune_neurons, utils/prune.py:950-977). */
int fb_forward_f64(const fb_model *m, int64_t rows, const double *X, const uint32_t *dead_bits, double *logits, void *stream);
int fb_forward_f32(const fb_model *m, int64_t rows, const float *X, const uint32_t *dead_bits, float *logits, void *stream);

/* -- pair screen: the predictions of CausalDiscriminationDetector.causal_discrimination, src/AC/metrics.py:130-159.
 *    X float32 [rows][d]; pa_cols (host) = protected columns; alts float32 [k][n_pa] (device) = every protected
 *    assignment in itertools.product order.  orig_pred[r] = class of row r; first_flip[r] = index of the first
 *    alternative (different from the row's own assignment) whose class differs, or -1. */
int fb_pair_screen(const fb_model *m, int64_t rows, const float *X, int n_pa, const int32_t *pa_cols_host, int k_alts,
                   const float *alts, uint8_t *orig_pred, int32_t *first_flip, void *stream);

/* -- per-partition sampled counterexample screen (the property src/AC/Verify-AC.py:137-156 gives Z3).
 *    ce_sample/pa_pos/pa_neg int32 [n]: lowest sample index that is a counterexample and the protected
 *    values with logit > 0 / < 0, or -1.  verdict (may be NULL) gets FB_VERDICT_CE OR-ed in.
 *    (fb_partition_screen_ex below with ra_col = -1 and no record.) */
int fb_partition_screen(const fb_model *m, int64_t n, const int32_t *boxes, const int32_t *samples, int S,
                        uint64_t seed, int64_t first_index, int pa_col, int32_t *ce_sample, int32_t *pa_pos,
                        int32_t *pa_neg, uint8_t *verdict, void *stream);

/* Same screen with every option.
 *    Limits: the protected column's range hi - lo + 1 of a box may be at most 65 536 values (256 with a relaxed column);
 *    a box beyond that is skipped and marked FB_VERDICT_SKIPPED in verdict / record.
 *    ra_col >= 0 selects the RELAXED query of relaxed/AC/Verify-AC.py:48-51, :128-132 (in_const_diff_adult,
 *    utils/verif_utils.py:859-867): the two points may also differ in column ra_col by at most ra_threshold (0..127).
 *    As in the reference, only the first point must lie in the box (in_const_domain_* bounds x' on the protected columns
 *    alone, :846-854): the second point's ra value ranges over a - thr .. a + thr around the sample's value a.
 *    A strict pair (difference 0) is preferred; otherwise the first second point -- difference ascending from -thr,
 *    protected value ascending -- with a partner of another protected value and the other class.
 *    ra_dpos / ra_dneg int32 [n] (may be NULL): the difference applied to the positive / negative point (0 for the point
 *    inside the box).  Any of ce_sample / pa_pos / pa_neg / verdict may be NULL.  record (may be NULL): the packed records
 *    fb_ibp wrote ([n][2*NW+2] uint32): the last word of each receives ce_sample, word 0 the verdict flags.
 *    skip (uint8 [n], may be NULL): boxes with a non-zero byte are not searched and report no counterexample.  Meant for
 *    boxes fb_ibp certified (FB_VERDICT_CERTIFIED: the logit is one-signed over the whole box, so no pair of points of
 *    the box can differ in class): for a strict query pass the certified bit of the boxes themselves, for a relaxed one
 *    that of the boxes widened by ra_threshold on both sides of column ra_col (the second point may leave the box
 *    there).  The certificate is about the real-valued network; the screen evaluates in float32. */
int fb_partition_screen_ex(const fb_model *m, int64_t n, const int32_t *boxes, const int32_t *samples, int S,
                           uint64_t seed, int64_t first_index, int pa_col, int ra_col, int ra_threshold,
                           int32_t *ce_sample, int32_t *pa_pos, int32_t *pa_neg, int32_t *ra_dpos, int32_t *ra_dneg,
                           uint8_t *verdict, uint32_t *record, const uint8_t *skip, void *stream);

/* -- hybrid prediction: find_partition_result_for_point + hybrid_predict, src/AC/Verify-AC-experiment-new2.py:587-638.
 *    For each of `rows` points: the partition of grid g that contains it, that partition's recorded verdict
 *    status[index] (uint8 [grid size]: FB_STATUS_NONE = not evaluated, _SAT, _UNSAT, _UNKNOWN), and the float32 logit
 *    of the model the verdict selects: `fairer` where the partition is unfair (SAT), `original` everywhere else (also
 *    for points outside every partition).  X64 float64 [rows][d] (located) and X32 float32 [rows][d] (predicted) hold the
 *    same points.  logit float32 [rows]; route uint8 [rows] (may be NULL) = the status that decided; counters uint32 [4]
 *    (may be NULL) += number of rows per status (the reference's debug_counters). */
#define FB_STATUS_NONE 0
#define FB_STATUS_SAT 1
#define FB_STATUS_UNSAT 2
#define FB_STATUS_UNKNOWN 3
int fb_hybrid_predict(const fb_grid *g, const fb_model *original, const fb_model *fairer, int64_t rows, const double *X64,
                      const float *X32, const uint8_t *status, float *logit, uint8_t *route, uint32_t *counters, void *stream);

/* -- diagnostic: how close the sampled pre-activations come to zero.  The reference's forward pass sums through numpy's
 *    BLAS in an unspecified order (utils/AC-1-Model-Functions.py:19); every kernel here sums in ascending input order.
 *    A sign -- all the activation profile consumes -- can differ between the two only where |pre-activation| is within
 *    rounding error of sum |w x| + |b|.  count (device uint64) = number of (sample, neuron) pairs of the batch with
 *    |pre| <= 2^-40 (sum |w x| + |b|); min_ratio (device double) = the smallest such ratio seen.  Not a fast path. */
int fb_profile_nearzero(const fb_model *m, int64_t n, const int32_t *boxes, const int32_t *samples, int S, uint64_t seed,
                        int64_t first_index, uint64_t *count, double *min_ratio, void *stream);

/* -- heuristic_prune, utils/prune.py:862-939 (only reached after a Z3 `unknown`): per hidden layer compare the
 *    ws_ub of sample-candidates and non-candidates (numpy mean / median / percentile semantics) and mark small-bound
 *    candidates dead.  ws_lb, ws_ub [n][N]; cand = `candidates`, s_cand = `s_candidates`, deads = `deads`, all [n][NW];
 *    pct = perc_threshold.  Outputs new_deads and merged_deads [n][NW], keep-one fix applied to both. */
int fb_heuristic(const fb_model *m, int64_t n, const double *ws_lb, const double *ws_ub, const uint32_t *cand,
                 const uint32_t *s_cand, const uint32_t *deads, double pct, uint32_t *new_deads, uint32_t *merged,
                 void *stream);

/* -- measurement aid: sustained DFMA issue rate of the current device, in TFLOP/s (FMA = 2 flops).
 *    This is the roofline denominator of the profile / interval kernels. */
int fb_fp64_peak(int iters, double *tflops, double *ms);

#ifdef __cplusplus
}
#endif
#endif /* FAIRIFY_B200_H */
