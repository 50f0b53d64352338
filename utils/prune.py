"""Drop-in for the reference's utils/prune.py, with the pre-solver stage on the GPU.

Same function names, positional signatures and return structures as the reference (plain lists of lists, a
pandas `sim_df`), so `from utils.prune import *` keeps the unchanged drivers working:

    bounds, candidates, s_candidates, b_deads, s_deads, st_deads, pos_prob, sim_X_df = \
        sound_prune(df, w, b, simulation_size, layer_net, p)              # src/AC/Verify-AC.py:117-118

What runs where
  * sampling keeps the reference's RNG: `np.random.randint(lo, hi+1, size=(S, d))` consumes the global legacy
    stream exactly like the scalar loop of `simluate_data` (utils/prune.py:205-222) -- same seed, same samples;
  * activation profile, interval bounds, masks, singular verification, keep-one fixes, merge: one `fb_prune` call on
    the device (`candidate_dead_nodes` :168-192, `neuron_bounds` :105-164, `dead_node_from_bound` :226-251,
    `singular_verification*` :276-644, :689-705);
  * the per-neuron Z3 query of `singular_verification*` is answered on the device from the closed form of what the
    reference asks (for a hidden layer l >= 1 it boxes the layer's inputs with the layer's OWN post-activation bounds,
    :266-273 called with layer_index at :302-345 -- see include/fairify_b200.h, fb_ibp).  Only neurons whose
    float result is within rounding error of zero ("slivers") go to Z3 on the host.  FAIRIFY_GPU_SINGULAR=reference
    runs the host query for every remaining candidate like the reference, =off never runs it;
  * batching behind the unbatched API: when `range_dict` is a PartitionBox of a PartitionList
    (utils/input_partition.py), the call computes a whole WINDOW of the list -- the next FAIRIFY_GPU_WINDOW (4096)
    partitions in the list's current order, so `shuffle(p_list)` is honoured -- in one launch and serves the following
    calls from it.  The samples of the window are drawn partition by partition from the global numpy stream and the
    stream state is put back to where the reference would have left it after every served call; a window is dropped
    as soon as the caller does anything the window did not foresee (another partition order, another model, random
    draws of its own in between), so results and RNG consumption are those of the one-at-a-time path
    (FAIRIFY_GPU_WINDOW=1);
  * `heuristic_prune` (:862-939) runs on the device too (`fb_heuristic`);
  * `prune_neurons`, `compression_ratio`, `merge_dead_nodes` operate on a handful of Python lists per
    partition and feed Z3; they stay host code.
"""
import copy
import os

import numpy as np
import pandas as pd

from utils import _device
from utils.verif_utils import *  # noqa: F401,F403  (the reference's prune.py leaks these names to the drivers too)

LABELS = {'adult': 'income-per-year', 'bank': 'y', 'german': 'credit', 'compass': 'score_factor',
          'default': 'default.payment.next.month'}


# ----------------------------------------------------------------------------------------------------
# small host helpers (list-of-lists <-> flat arrays)
# ----------------------------------------------------------------------------------------------------
def _split_lists(model, flat, cast):
    return [[cast(v) for v in part] for part in model.split(np.asarray(flat))]


def _box_from(cols, range_dict):
    return np.array([[range_dict[c][0], range_dict[c][1]] for c in cols], dtype=np.int32)


def _keep_one(mask):
    for l in mask:
        if 0 not in l:
            l[0] = 0
    return mask


# ----------------------------------------------------------------------------------------------------
# reference API, piece by piece
# ----------------------------------------------------------------------------------------------------
def simluate_data(df, size, range_dict):
    """`size` uniform integer rows inside the box (reference :205-222), drawn from the global numpy stream in
    the same order as the reference's scalar loop."""
    cols = df.columns.values
    lo = np.array([range_dict[c][0] for c in cols]); hi = np.array([range_dict[c][1] for c in cols])
    sim = np.random.randint(lo, hi + 1, size=(size, len(cols)))
    return pd.DataFrame(data=sim, columns=cols)


def candidate_dead_nodes(data, weight, bias, layer_net):
    """(dead mask, positive_prob): neurons that are never non-zero on the samples (reference :168-192)."""
    m = _device.model_for(weight, bias)
    data = np.asarray(data)
    samples = np.ascontiguousarray(data, dtype=np.int32)[None]
    if not np.array_equal(samples[0], data):
        raise TypeError("candidate_dead_nodes expects integer sample rows (it gets simluate_data's output)")
    counts = m.profile(np.zeros((1, m.d, 2), dtype=np.int32), samples=samples).cpu().numpy()[0]
    dead = _split_lists(m, (counts == 0).astype(np.int64), int)
    prob = _split_lists(m, counts / data.shape[0], float)
    return dead, prob


def neuron_bounds(df, w, b, range_dict):
    """(ws_lb, ws_ub, pl_lb, pl_ub) exactly as the reference's float64 recurrence computes them (:105-164)."""
    m = _device.model_for(w, b)
    box = _box_from(df.columns.values, range_dict)[None]
    r = m.ibp(box, want=("ws_lb", "ws_ub", "pl_lb", "pl_ub"))
    return tuple(_split_lists(m, r.numpy(k)[0], float) for k in ("ws_lb", "ws_ub", "pl_lb", "pl_ub"))


def dead_node_from_bound(cand, weight, bias, range_dict, ws_ub):
    """Candidates whose upper bound is <= 0 are dead; last layer untouched (reference :226-251)."""
    print('INTERVAL BASED PRUNING')
    candidates = copy.deepcopy(cand)
    dead = copy.deepcopy(cand)
    finds = total = 0
    for l in range(len(ws_ub) - 1):
        for j in range(len(ws_ub[l])):
            total += 1
            if candidates[l][j] == 0:
                continue
            if ws_ub[l][j] <= 0:
                finds += 1; dead[l][j] = 1; candidates[l][j] = 0
            else:
                dead[l][j] = 0
    return dead, candidates, finds / total


def compression_ratio(deads):
    total = sum(len(l) for l in deads)
    alive = sum(1 for l in deads for d in l if not d)
    return 1 - (alive / total)


def merge_dead_nodes(b, s):
    merged = copy.deepcopy(b)
    for l in range(len(s)):
        for j in range(len(s[l])):
            if s[l][j] == 1:
                merged[l][j] = 1
    return merged


def prune_neurons(weight, bias, candidates):
    """Delete dead columns / the next layer's rows (reference :950-977); feeds the Z3 encoder on the host.
    The result is remembered as "these weights with that mask", so that predictions with the pruned network
    (`get_y_pred(net, pr_w, pr_b, ...)`, the C-check) run on the resident original model with the dead neurons
    masked instead of uploading a new model per partition."""
    pr_w, pr_b = [], []
    keep_prev = None
    for i in range(len(weight)):
        keep = ~np.asarray(candidates[i], dtype=bool)
        wi = np.asarray(weight[i])
        if keep_prev is not None:
            wi = wi[keep_prev, :]
        pr_w.append(wi[:, keep]); pr_b.append(np.asarray(bias[i])[keep])
        keep_prev = keep
    _device.remember_pruned(pr_w, pr_b, weight, bias, candidates)
    print('Pruning done!')
    return pr_w, pr_b


# ----------------------------------------------------------------------------------------------------
# singular verification (host, Z3)
# ----------------------------------------------------------------------------------------------------
def _z3_positive_is_unsat(l, j, cols, weight, bias, ranges, pl_lb, pl_ub):
    """The reference's query for one neuron: input-domain constraints + y_j > 0; True when unsat.
    Layer 0 ranges over Int variables inside the box (:298-300); deeper layers over Real variables bounded by
    pl_lb[l][k] <= x_k <= pl_ub[l][k] for k < len(pl_lb[l]) -- the reference indexes the post-activation
    bounds with the layer's own index (:302-305, :266-273) and this restatement keeps that."""
    import z3
    W = np.asarray(weight[l], dtype=np.float64).astype(object)
    B = np.asarray(bias[l], dtype=np.float64).astype(object)
    if l == 0:
        xi = [z3.Int('x%s' % i) for i in range(W.shape[0])]
        props = [z3.And(xi[i] >= ranges[c][0], xi[i] <= ranges[c][1]) for i, c in enumerate(cols)]
        x = np.array([z3.ToReal(v) for v in xi], dtype=object)
    else:
        x = np.array([z3.Real('x%s' % i) for i in range(W.shape[0])], dtype=object)
        props = []
        for k in range(len(pl_lb[l])):
            props.append(x[k] >= pl_lb[l][k]); props.append(x[k] <= pl_ub[l][k])
    y = W.T @ x + B
    s = z3.Solver()
    for p in props:
        s.add(p)
    s.add(y[j] > 0)
    return s.check() == z3.unsat


def _singular(cand, df, weight, bias, ranges, pl_lb, pl_ub, only=None, otherwise=None):
    """The reference's loop (:276-356).  `only` restricts the Z3 query to the flagged neurons; the others take
    their answer from `otherwise` (the device's closed form)."""
    print('SINGULAR VERIFICATION')
    candidates = copy.deepcopy(cand)
    dead = copy.deepcopy(cand)
    cols = list(df.columns.values)
    finds = total = 0
    for l in range(len(candidates) - 1):
        if l >= 1 and any(candidates[l]) and len(candidates[l]) > np.asarray(weight[l]).shape[0]:
            raise IndexError("index %d is out of bounds for axis 0 with size %d" %
                             (np.asarray(weight[l]).shape[0], np.asarray(weight[l]).shape[0]))   # as the reference (:269-270)
        for j in range(len(candidates[l])):
            total += 1
            if candidates[l][j] == 0:
                continue
            if only is None or only[l][j]:
                unsat = _z3_positive_is_unsat(l, j, cols, weight, bias, ranges, pl_lb, pl_ub)
            else:
                unsat = bool(otherwise[l][j]) if otherwise is not None else False
            if unsat:
                finds += 1; dead[l][j] = 1; candidates[l][j] = 0
            else:
                dead[l][j] = 0
    return dead, candidates, (finds / total if total else 0.0)


def singular_verification(cand, df, weight, bias, ranges, pl_lb, pl_ub):
    return _singular(cand, df, weight, bias, ranges, pl_lb, pl_ub)


singular_verification_german = singular_verification_bank = singular_verification
singular_verification_compass = singular_verification_default = singular_verification


# ----------------------------------------------------------------------------------------------------
# sound_prune*: the drop-in entry points (reference :671-859; they differ only in the label column)
# ----------------------------------------------------------------------------------------------------
_WANT = ("cand", "b_dead", "dead", "sliver", "s_dead", "s_cand", "counts", "ws_lb", "ws_ub", "pl_lb", "pl_ub")


class _Window:
    """Results of one batched launch: `items[k]` is the PartitionBox served k-th."""

    def __init__(self, key, items, sims, states, arrays):
        self.key, self.items, self.sims, self.states, self.arrays = key, items, sims, states, arrays
        self.next = 0


_window = None


def _same_rng_state(a, b):
    return a[2] == b[2] and a[3] == b[3] and a[4] == b[4] and np.array_equal(a[1], b[1])


def _window_size():
    try:
        return max(1, int(os.environ.get("FAIRIFY_GPU_WINDOW", "4096")))
    except ValueError:
        return 4096


def _compute(m, cols, boxes, S):
    """Draw the samples of `boxes` (a list of {attr: [lo, hi]}) one partition after the other from the global numpy
    stream, run the stage for all of them in one launch, bring everything the 8-tuple needs to the host."""
    n, d = len(boxes), len(cols)
    box = np.empty((n, d, 2), dtype=np.int32)
    sims = np.empty((n, S, d), dtype=np.int64)
    states = []
    for k, rd in enumerate(boxes):
        box[k] = _box_from(cols, rd)
        sims[k] = np.random.randint(box[k, :, 0].astype(np.int64), box[k, :, 1].astype(np.int64) + 1, size=(S, d))
        states.append(np.random.get_state())
    lo, hi = box[:, :, 0].min(axis=0), box[:, :, 1].max(axis=0)
    tol = m.tolerance(lo, hi)
    r = m.prune(box, samples=sims.astype(np.int32), tol=tol, want=_WANT)
    arrays = {k: r.numpy(k) for k in ("ws_lb", "ws_ub", "pl_lb", "pl_ub")}
    arrays["counts"] = r.numpy("counts").view(np.uint16)
    for k in ("cand", "b_dead", "dead", "sliver", "s_dead", "s_cand"):
        arrays[k] = r.bits(k)
    return sims, states, arrays


_xdf_cache = {}


def _feature_frame(df, label_name):
    """df without its label column (the reference drops it on every call, utils/prune.py:674; the drivers pass the
    same frame for every partition, so the result is kept while the frame's column index object is the same)."""
    hit = _xdf_cache.get(id(df))
    if hit is not None and hit[0] is df.columns and hit[1] == label_name:
        return hit[2]
    x_df = df.drop(labels=[label_name], axis=1, inplace=False)
    _xdf_cache.clear()
    _xdf_cache[id(df)] = (df.columns, label_name, x_df, df)       # (df kept alive: its id cannot be reused)
    return x_df


def _sound_prune(label_name, df, weight, bias, simulation_size, layer_net, range_dict):
    global _window
    x_df = _feature_frame(df, label_name)
    cols = x_df.columns.values
    m = _device.model_for(weight, bias)
    S = int(simulation_size)
    key = (id(m), S, tuple(cols))
    plist = getattr(range_dict, "grid_ref", None)
    w = _window
    # the window serves this call only if it is the call the window expects next AND nobody else has drawn from numpy's
    # global stream since the previous call (the window's samples were drawn assuming nobody would)
    if not (w is not None and w.key == key and w.next < len(w.items) and w.items[w.next] is range_dict
            and _same_rng_state(np.random.get_state(), w.states[w.next - 1])):
        # a new window: the next partitions of the list in its current order, starting at this one
        items = [range_dict]
        wsize = _window_size()
        if plist is not None and wsize > 1:
            pos = plist.position(range_dict)
            if pos is not None:
                items = plist[pos:pos + wsize]
        sims, states, arrays = _compute(m, cols, items, S)
        w = _window = _Window(key, items, sims, states, arrays)
    k = w.next
    w.next += 1
    np.random.set_state(w.states[k])           # the global stream is where the reference would have left it
    a = {name: arr[k] for name, arr in w.arrays.items()}
    sim_df = pd.DataFrame(data=w.sims[k], columns=cols)
    if w.next >= len(w.items):
        _window = None

    bounds = tuple(_split_lists(m, a[name], float) for name in ("ws_lb", "ws_ub", "pl_lb", "pl_ub"))
    candidates = _split_lists(m, a["cand"], int)
    pos_prob = _split_lists(m, a["counts"] / simulation_size, float)
    b_dead = _split_lists(m, a["b_dead"], int)
    print('INTERVAL BASED PRUNING')
    mode = os.environ.get("FAIRIFY_GPU_SINGULAR", "device")
    if mode == "device" and not a["sliver"].any():
        print('SINGULAR VERIFICATION')
        s_cand = _split_lists(m, a["s_cand"], int)
        s_dead = _split_lists(m, a["s_dead"], int)
        dead_nodes = _split_lists(m, a["dead"], int)
        return bounds, candidates, s_cand, b_dead, s_dead, dead_nodes, pos_prob, sim_df

    # b_candidates: candidates not proven dead by the bound; last layer copied through (:235-251)
    le0 = a["ws_ub"] <= 0
    b_cand_flat = a["cand"] & ~le0
    b_cand_flat[m.boff[m.L - 1]:] = a["cand"][m.boff[m.L - 1]:]
    b_cand = _split_lists(m, b_cand_flat, int)
    if mode == "reference":
        s_dead, s_cand, _ = _singular(b_cand, x_df, weight, bias, range_dict, bounds[2], bounds[3])
    elif mode == "off":
        print('SINGULAR VERIFICATION')
        s_cand = copy.deepcopy(b_cand)
        s_dead = [[0] * len(l) for l in b_cand]
        s_dead[-1] = list(b_cand[-1])           # the reference copies the last layer through (:287-288)
    else:
        # the device's answer everywhere except on the slivers, which get the reference's exact query
        sliver = _split_lists(m, a["sliver"], int)
        raw_dead = _split_lists(m, b_cand_flat & ~a["s_cand"], int)       # s_dead before its keep-one fix
        s_dead, s_cand, _ = _singular(b_cand, x_df, weight, bias, range_dict, bounds[2], bounds[3], only=sliver,
                                      otherwise=raw_dead)
    _keep_one(s_dead)
    dead_nodes = _keep_one(merge_dead_nodes(b_dead, s_dead))
    return bounds, candidates, s_cand, b_dead, s_dead, dead_nodes, pos_prob, sim_df


def sound_prune(df, weight, bias, simulation_size, layer_net, range_dict):
    return _sound_prune(LABELS['adult'], df, weight, bias, simulation_size, layer_net, range_dict)


def sound_prune_bank(df, weight, bias, simulation_size, layer_net, range_dict):
    return _sound_prune(LABELS['bank'], df, weight, bias, simulation_size, layer_net, range_dict)


def sound_prune_german(df, weight, bias, simulation_size, layer_net, range_dict):
    return _sound_prune(LABELS['german'], df, weight, bias, simulation_size, layer_net, range_dict)


def sound_prune_compass(df, weight, bias, simulation_size, layer_net, range_dict):
    return _sound_prune(LABELS['compass'], df, weight, bias, simulation_size, layer_net, range_dict)


def sound_prune_default(df, weight, bias, simulation_size, layer_net, range_dict):
    return _sound_prune(LABELS['default'], df, weight, bias, simulation_size, layer_net, range_dict)


def sound_prune_global(df, weight, bias, simulation_size, layer_net, range_dict):
    """Reference :646-667: label = last column, returns the pruned weights."""
    out = _sound_prune(df.columns[len(df.columns) - 1], df, weight, bias, simulation_size, layer_net, range_dict)
    return prune_neurons(weight, bias, out[5])


def harsh_prune(df, weight, bias, simulation_size, layer_net, range_dict):
    """Reference :89-102: prune every sample-candidate (unsound), label = last column."""
    x_data = df.drop(df.columns[len(df.columns) - 1], axis=1, inplace=False)
    sim_df = simluate_data(x_data, simulation_size, range_dict)
    candidates, _ = candidate_dead_nodes(sim_df.to_numpy(), weight, bias, layer_net)
    print(candidates)
    pr_w, pr_b = prune_neurons(weight, bias, candidates)
    print(round(compression_ratio(candidates), 2), '% pruning')
    return pr_w, pr_b


# ----------------------------------------------------------------------------------------------------
# heuristic_prune (reference :862-939): only reached after a Z3 `unknown`; a few numpy statistics per layer
# ----------------------------------------------------------------------------------------------------
def heuristic_prune(bounds, candidates, s_candidates, deads, pos_prob, perc_threshold, w, b):
    """(new_deads, merged_deads), computed on the device (`fb_heuristic`) from the reference's list arguments."""
    from fairify_b200.engine import pack_bits, unpack_bits
    ws_lb, ws_ub, pl_lb, pl_ub = bounds
    m = _device.model_for(w, b)
    flat = lambda ll: np.concatenate([np.asarray(l, dtype=np.float64) for l in ll])[None]
    bits = lambda ll: pack_bits(np.concatenate([np.asarray(l, dtype=np.uint8) != 0 for l in ll]).astype(np.uint8)[None])
    new, merged = m.heuristic(flat(ws_lb), flat(ws_ub), bits(candidates), bits(s_candidates), bits(deads), perc_threshold)
    new = unpack_bits(new.cpu().numpy().view(np.uint32), m.N)[0]
    merged = unpack_bits(merged.cpu().numpy().view(np.uint32), m.N)[0]
    return _split_lists(m, new, int), _split_lists(m, merged, int)
