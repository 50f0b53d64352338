"""ctypes front end of the CPU oracle (oracle/fairify_oracle.c).

TEST INFRASTRUCTURE ONLY -- see the header of fairify_oracle.c.  Nothing under fairify_b200/ or utils/
imports this module; only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs do.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libfairify_oracle.so")
_lib = None


def build(force=False):
    src = os.path.join(_HERE, "fairify_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return _SO


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = C.CDLL(_SO)
        _lib.fo_grid_count.restype = C.c_int64
        _lib.fo_compression_ratio.restype = C.c_double
        for f in ("fo_np_sum", "fo_np_mean", "fo_np_median_sorted", "fo_np_percentile_sorted"):
            getattr(_lib, f).restype = C.c_double
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class Net:
    """Weights packed the way the oracle (and the device library) take them."""

    def __init__(self, w, b):
        self.L = len(w)
        self.dims = np.array([w[0].shape[0]] + [x.shape[1] for x in w], dtype=np.int32)
        self.W32 = np.concatenate([np.ascontiguousarray(x, dtype=np.float32).ravel() for x in w])
        self.B32 = np.concatenate([np.ascontiguousarray(x, dtype=np.float32).ravel() for x in b])
        self.W = self.W32.astype(np.float64)
        self.B = self.B32.astype(np.float64)
        self.N = int(self.dims[1:].sum())
        self.P = int(self.W.size)
        self.d = int(self.dims[0])
        self.boff = np.concatenate([[0], np.cumsum(self.dims[1:])]).astype(np.int64)

    def split(self, flat):
        """[N] (or [n, N]) flat per-neuron array -> list of per-layer arrays."""
        return [flat[..., self.boff[l]:self.boff[l + 1]] for l in range(self.L)]


def grid_count(lo, hi, T):
    lo = np.ascontiguousarray(lo, dtype=np.int64); hi = np.ascontiguousarray(hi, dtype=np.int64)
    nch = np.zeros(len(lo), dtype=np.int64)
    n = lib().fo_grid_count(len(lo), _p(lo), _p(hi), C.c_int64(T), _p(nch))
    return int(n), nch


def grid_boxes(lo, hi, T, key_order, first, count):
    lo = np.ascontiguousarray(lo, dtype=np.int64); hi = np.ascontiguousarray(hi, dtype=np.int64)
    ko = np.ascontiguousarray(key_order, dtype=np.int32)
    out = np.zeros((count, len(lo), 2), dtype=np.int32)
    lib().fo_grid_boxes(len(lo), _p(lo), _p(hi), C.c_int64(T), _p(ko), C.c_int64(first), C.c_int64(count), _p(out))
    return out


def grid_locate(lo, hi, T, key_order, X):
    lo = np.ascontiguousarray(lo, dtype=np.int64); hi = np.ascontiguousarray(hi, dtype=np.int64)
    ko = np.ascontiguousarray(key_order, dtype=np.int32)
    X = np.ascontiguousarray(X, dtype=np.float64); m = X.shape[0]
    out = np.zeros(m, dtype=np.int64)
    lib().fo_grid_locate(len(lo), _p(lo), _p(hi), C.c_int64(T), _p(ko), C.c_int64(m), _p(X), _p(out))
    return out


def ibp(net, boxes):
    boxes = np.ascontiguousarray(boxes, dtype=np.int32); n = boxes.shape[0]
    outs = [np.zeros((n, net.N)) for _ in range(4)]
    lib().fo_ibp(net.L, _p(net.dims), _p(net.W), _p(net.B), C.c_int64(n), _p(boxes), *[_p(o) for o in outs])
    return outs


def ibp_directed(net, boxes):
    boxes = np.ascontiguousarray(boxes, dtype=np.int32); n = boxes.shape[0]
    outs = [np.zeros((n, net.N)) for _ in range(2)]
    lib().fo_ibp_directed(net.L, _p(net.dims), _p(net.W), _p(net.B), C.c_int64(n), _p(boxes), *[_p(o) for o in outs])
    return outs


def philox_samples(seed, first_index, boxes, S):
    boxes = np.ascontiguousarray(boxes, dtype=np.int32); n, d = boxes.shape[0], boxes.shape[1]
    out = np.zeros((n, S, d), dtype=np.int32)
    lib().fo_philox_samples(C.c_uint64(seed), C.c_int64(first_index), C.c_int64(n), S, d, _p(boxes), _p(out))
    return out


def profile(net, samples):
    samples = np.ascontiguousarray(samples, dtype=np.int32); n, S = samples.shape[0], samples.shape[1]
    counts = np.zeros((n, net.N), dtype=np.uint32)
    lib().fo_profile(net.L, _p(net.dims), _p(net.W), _p(net.B), C.c_int64(n), S, _p(samples), _p(counts))
    return counts


def forward_f64(net, X):
    X = np.ascontiguousarray(X, dtype=np.float64); m = X.shape[0]
    out = np.zeros(m)
    lib().fo_forward_f64(net.L, _p(net.dims), _p(net.W), _p(net.B), C.c_int64(m), _p(X), _p(out))
    return out


def forward_f32(net, X):
    X = np.ascontiguousarray(X, dtype=np.float32); m = X.shape[0]
    out = np.zeros(m, dtype=np.float32)
    lib().fo_forward_f32(net.L, _p(net.dims), _p(net.W32), _p(net.B32), C.c_int64(m), _p(X), _p(out))
    return out


def masks(net, counts, ws_ub, pl_lb, pl_ub, tol=None):
    """Mask algebra of sound_prune* for [n, N] inputs (counts may be None: every neuron a candidate).  Returns byte
    masks cand, b_dead, b_cand, s_dead, s_cand, dead, sliver and `q`, the float value of the quantity the
    reference's Z3 neuron query decides the sign of (see singular_one in fairify_oracle.c)."""
    ws_ub = np.ascontiguousarray(ws_ub, dtype=np.float64); n = ws_ub.shape[0]
    if counts is not None:
        counts = np.ascontiguousarray(counts, dtype=np.uint32)
    pl_lb = np.ascontiguousarray(pl_lb, dtype=np.float64); pl_ub = np.ascontiguousarray(pl_ub, dtype=np.float64)
    if tol is not None:
        tol = np.ascontiguousarray(tol, dtype=np.float64)
    names = ("cand", "b_dead", "b_cand", "s_dead", "s_cand", "dead", "sliver")
    outs = {k: np.zeros((n, net.N), dtype=np.uint8) for k in names}
    q = np.zeros((n, net.N))
    lib().fo_masks(net.L, _p(net.dims), _p(net.W), _p(net.B), C.c_int64(n), _p(counts) if counts is not None else None,
                   _p(ws_ub), _p(pl_lb), _p(pl_ub), _p(tol) if tol is not None else None,
                   *[_p(outs[k]) for k in names], _p(q))
    outs["q"] = q
    return outs


def compression_ratio(mask):
    mask = np.ascontiguousarray(mask, dtype=np.uint8)
    return lib().fo_compression_ratio(int(mask.size), _p(mask))


def heuristic_prune(net, ws_lb, ws_ub, cand, s_cand, deads, pct):
    a = [np.ascontiguousarray(x, dtype=np.float64) for x in (ws_lb, ws_ub)]
    mk = [np.ascontiguousarray(x, dtype=np.uint8) for x in (cand, s_cand, deads)]
    new = np.zeros(net.N, dtype=np.uint8); merged = np.zeros(net.N, dtype=np.uint8)
    lib().fo_heuristic_prune(net.L, _p(net.dims), _p(a[0]), _p(a[1]), _p(mk[0]), _p(mk[1]), _p(mk[2]),
                             C.c_double(pct), _p(new), _p(merged))
    return new, merged


def pair_screen(net, X, pa_cols, alts):
    X = np.ascontiguousarray(X, dtype=np.float32); m = X.shape[0]
    pa = np.ascontiguousarray(pa_cols, dtype=np.int32)
    alts = np.ascontiguousarray(alts, dtype=np.float32).reshape(-1, len(pa))
    orig = np.zeros(m, dtype=np.uint8); flip = np.zeros(m, dtype=np.int32)
    lib().fo_pair_screen(net.L, _p(net.dims), _p(net.W32), _p(net.B32), C.c_int64(m), _p(X), len(pa), _p(pa),
                         alts.shape[0], _p(alts), _p(orig), _p(flip))
    return orig, flip


def partition_screen(net, samples, boxes, pa_col):
    samples = np.ascontiguousarray(samples, dtype=np.int32); boxes = np.ascontiguousarray(boxes, dtype=np.int32)
    n, S = samples.shape[0], samples.shape[1]
    outs = [np.zeros(n, dtype=np.int32) for _ in range(3)]
    lib().fo_partition_screen(net.L, _p(net.dims), _p(net.W32), _p(net.B32), C.c_int64(n), S, _p(samples),
                              _p(boxes), int(pa_col), *[_p(o) for o in outs])
    return outs


def partition_screen_ex(net, samples, boxes, pa_col, ra_col=-1, thr=0):
    """(ce_sample, pa_pos, pa_neg, ra_dpos, ra_dneg): the screen with the relaxed query (fo_partition_screen_ex)."""
    samples = np.ascontiguousarray(samples, dtype=np.int32); boxes = np.ascontiguousarray(boxes, dtype=np.int32)
    n, S = samples.shape[0], samples.shape[1]
    outs = [np.zeros(n, dtype=np.int32) for _ in range(5)]
    lib().fo_partition_screen_ex(net.L, _p(net.dims), _p(net.W32), _p(net.B32), C.c_int64(n), S, _p(samples), _p(boxes),
                                 int(pa_col), int(ra_col), int(thr), *[_p(o) for o in outs])
    return outs


def nearzero(net, samples):
    """(count of (sample, neuron) pairs with |pre| <= 2^-40 (sum |w x| + |b|), smallest ratio)."""
    samples = np.ascontiguousarray(samples, dtype=np.int32); n, S = samples.shape[0], samples.shape[1]
    cnt = C.c_uint64(); best = C.c_double()
    lib().fo_nearzero(net.L, _p(net.dims), _p(net.W), _p(net.B), C.c_int64(n), S, _p(samples), C.byref(cnt), C.byref(best))
    return int(cnt.value), float(best.value)


def hybrid_predict(original, fairer, lo, hi, T, key_order, X, status):
    """find_partition_result_for_point + hybrid_predict (src/AC/Verify-AC-experiment-new2.py:587-638) restated:
    (float32 logit of the selected model, route = status that decided) per row of X."""
    X = np.ascontiguousarray(X, dtype=np.float64)
    idx = grid_locate(lo, hi, T, key_order, X)
    st = np.where(idx >= 0, np.asarray(status, dtype=np.uint8)[np.maximum(idx, 0)], 0).astype(np.uint8)
    st = np.where(st <= 3, st, 0).astype(np.uint8)
    yo = forward_f32(original, X.astype(np.float32)); yf = forward_f32(fairer, X.astype(np.float32))
    return np.where(st == 1, yf, yo).astype(np.float32), st


def prune_batch(net, boxes, seed, first_index, S, tol):
    boxes = np.ascontiguousarray(boxes, dtype=np.int32); n = boxes.shape[0]
    tol = np.ascontiguousarray(tol, dtype=np.float64)
    counts = np.zeros((n, net.N), dtype=np.uint32)
    names = ("cand", "b_dead", "s_dead", "dead", "sliver")
    outs = {k: np.zeros((n, net.N), dtype=np.uint8) for k in names}
    verdict = np.zeros(n, dtype=np.uint8)
    lib().fo_prune_batch(net.L, _p(net.dims), _p(net.W), _p(net.B), C.c_int64(n), _p(boxes), C.c_uint64(seed),
                         C.c_int64(first_index), S, _p(tol), _p(counts), *[_p(outs[k]) for k in names], _p(verdict))
    outs["counts"] = counts; outs["verdict"] = verdict
    return outs


def num_threads():
    return int(lib().fo_num_threads())


def set_threads(n):
    """Use n OpenMP threads from now on (torchrun exports OMP_NUM_THREADS=1, which would hide the host's cores)."""
    lib().fo_set_threads(int(n))


def sliver_tolerance(net, lo, hi):
    """Per-neuron tolerance below which the float sign of the singular-verification quantity is not trusted, valid
    for every box inside [lo, hi].

    Layer 0: the quantity is ws_ub; (in+2) roundings of relative size 2^-53 on terms whose magnitudes sum to at most
    M_j = sum_i |w_ij| * max|x_i| + |b_j|.  Layer l >= 1: the quantity is q_j (singular_one in fairify_oracle.c), a sum
    over the first min(in, out) inputs bounded by the layer's OWN post-activation magnitudes M_l, so the terms sum to at
    most sum_k |w_kj| M_l[k] + |b_j|; Z3 additionally reads every float through its decimal repr (one more relative
    2^-53 per factor).  Both with a x4 safety factor.  The same formula is used by the host library
    (fairify_b200.engine.sliver_tolerance); kept separately here so the oracle stays self-contained.
    """
    mag = np.maximum(np.abs(np.asarray(lo, dtype=np.float64)), np.abs(np.asarray(hi, dtype=np.float64)))
    tol = np.zeros(net.N)
    wo = 0
    for l in range(net.L):
        i, o = int(net.dims[l]), int(net.dims[l + 1])
        w = np.abs(net.W[wo:wo + i * o].reshape(i, o)); wo += i * o
        b = np.abs(net.B[net.boff[l]:net.boff[l + 1]])
        M = mag @ w + b
        if l == 0:
            t = 4.0 * (i + 2) * 2.0 ** -53 * M
        else:
            mm = min(i, o)
            t = 4.0 * (mm + 4) * 2.0 ** -53 * (M[:mm] @ w[:mm] + b)
        tol[net.boff[l]:net.boff[l + 1]] = t
        mag = M
    return tol
