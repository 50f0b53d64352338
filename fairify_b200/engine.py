"""Host API of the B200 pre-solver stage: thin Python over the C ABI (include/fairify_b200.h).

torch is used for what it is good at here -- device memory, pinned host memory, streams -- and nothing
else: every computation below is a call into libfairify_b200.so with raw device pointers.

Reference functions behind each method (paths relative to the Fairify checkout):
  Grid                 utils/input_partition.py:17-76   partition + partitioned_ranges
  Model.profile        utils/prune.py:168-192           candidate_dead_nodes (+ :205-222 simluate_data)
  Model.ibp            utils/prune.py:105-164, :226-251 neuron_bounds + dead_node_from_bound + keep-one fix
  Model.prune          utils/prune.py:671-859           sound_prune* minus the Z3 neuron queries
  Model.forward        utils/<M>-Model-Functions.py     net  (get_y_pred, utils/verif_utils.py:1040-1047)
  Model.pair_screen    src/AC/metrics.py:130-159        CausalDiscriminationDetector predictions
  Model.partition_screen  src/AC/Verify-AC.py:137-156   the fairness property, screened on samples
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import FairifyB200Error, PruneOut, check

MASK_OUTPUTS = ("cand", "b_dead", "dead", "sliver", "s_dead", "s_cand")
BOUND_OUTPUTS = ("ws_lb", "ws_ub", "pl_lb", "pl_ub", "lb_rd", "ub_ru")
DEFAULT_OUTPUTS = ("cand", "b_dead", "dead", "sliver", "verdict", "counts")
ALL_OUTPUTS = DEFAULT_OUTPUTS + BOUND_OUTPUTS + ("s_dead", "s_cand", "record")

VERDICT_CERTIFIED, VERDICT_CE, VERDICT_SLIVER = 1, 2, 4


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def _stream():
    """torch's current stream on the current device, as the C-ABI's cudaStream_t."""
    if _raw_stream is not None:                      # same value as below, without building a Stream object per call
        return C.c_void_p(_raw_stream(torch.cuda.current_device()))
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _require_cuda():
    if not torch.cuda.is_available():
        raise FairifyB200Error("no CUDA device: the pre-solver stage runs on the GPU only (there is no CPU fallback)")


def unpack_bits(words, N):
    """uint32 [n, NW] mask words (numpy) -> uint8 [n, N] of 0/1."""
    words = np.ascontiguousarray(words, dtype=np.uint32)
    bits = np.unpackbits(words.view(np.uint8).reshape(words.shape[0], -1), axis=1, bitorder="little")
    return bits[:, :N]


def pack_bits(bits):
    """uint8 [n, N] of 0/1 -> uint32 [n, NW]."""
    bits = np.asarray(bits, dtype=np.uint8)
    n, N = bits.shape
    NW = (N + 31) // 32
    padded = np.zeros((n, NW * 32), dtype=np.uint8)
    padded[:, :N] = bits
    return np.packbits(padded, axis=1, bitorder="little").view(np.uint32).reshape(n, NW)


def sliver_tolerance(w, b, lo, hi):
    """Per-neuron tolerance below which the float sign of the singular-verification quantity is not trusted, valid
    for every box inside [lo, hi] (argument `tol` of fb_ibp / fb_prune).

    The reference's Z3 query (singular_verification*, utils/prune.py:276-644) decides the sign of an exact sum; the
    device evaluates the same sum in float64.  Layer 0: the sum is ws_ub -- (in+2) roundings of relative size 2^-53 on
    terms whose magnitudes add up to at most M_j = sum_i |w_ij| max|x_i| + |b_j|.  Layer l >= 1: the sum is q_j, over
    the first min(in, out) inputs boxed by the layer's OWN post-activation bounds (see include/fairify_b200.h, fb_ibp),
    whose magnitudes are at most M_l, so its terms add up to at most sum_k |w_kj| M_l[k] + |b_j|; Z3 also reads every
    float through its decimal repr, one more relative 2^-53 per factor.  Both with a safety factor of 4.  A candidate
    whose |sum| is within the tolerance is a "sliver": the host puts the reference's own query to Z3 for it.
    """
    mag = np.maximum(np.abs(np.asarray(lo, dtype=np.float64)), np.abs(np.asarray(hi, dtype=np.float64)))
    out = []
    for l, (wl, bl) in enumerate(zip(w, b)):
        wl = np.abs(np.asarray(wl, dtype=np.float64)); bl = np.abs(np.asarray(bl, dtype=np.float64))
        M = mag @ wl + bl
        if l == 0:
            out.append(4.0 * (wl.shape[0] + 2) * 2.0 ** -53 * M)
        else:
            mm = min(wl.shape)
            out.append(4.0 * (mm + 4) * 2.0 ** -53 * (M[:mm] @ wl[:mm] + bl))
        mag = M
    return np.concatenate(out)


class PruneResult:
    """Device-resident outputs of one prune call (torch tensors), with host-side accessors."""

    def __init__(self, model, n, S, tensors):
        self.model, self.n, self.S, self.t = model, n, S, tensors

    def __getattr__(self, k):
        t = self.__dict__.get("t", {})
        if k in t:
            return t[k]
        raise AttributeError(k)

    def numpy(self, key):
        return self.t[key].cpu().numpy()

    def bits(self, key):
        """mask `key` as uint8 [n, N]."""
        return unpack_bits(self.numpy(key), self.model.N)


class Model:
    """A Dense/ReLU network with one logit, resident on the current CUDA device."""

    def __init__(self, w, b):
        _require_cuda()
        self.w = [np.ascontiguousarray(x, dtype=np.float32) for x in w]
        self.b = [np.ascontiguousarray(x, dtype=np.float32).ravel() for x in b]
        if len(self.w) != len(self.b) or not self.w:
            raise ValueError("weight and bias lists must have the same non-zero length")
        self.dims = np.array([self.w[0].shape[0]] + [x.shape[1] for x in self.w], dtype=np.int32)
        for l, (wl, bl) in enumerate(zip(self.w, self.b)):
            if wl.shape != (self.dims[l], self.dims[l + 1]) or bl.shape != (self.dims[l + 1],):
                raise ValueError(f"layer {l}: kernel {wl.shape} / bias {bl.shape} do not chain")
        self.L = len(self.w)
        self.d = int(self.dims[0])
        self.N = int(self.dims[1:].sum())
        self.P = int(sum(x.size for x in self.w))
        self.NW = (self.N + 31) // 32
        self.boff = np.concatenate([[0], np.cumsum(self.dims[1:])]).astype(np.int64)
        torch.cuda.current_stream()              # make sure torch's context exists on this device
        self.device = torch.device("cuda", torch.cuda.current_device())
        wflat = np.concatenate([x.ravel() for x in self.w]); bflat = np.concatenate(self.b)
        h = C.c_void_p()
        check(_lib.lib().fb_model_create(self.L, self.dims.ctypes.data_as(C.c_void_p), wflat.ctypes.data_as(C.c_void_p),
                                         bflat.ctypes.data_as(C.c_void_p), C.byref(h)), "fb_model_create")
        self._h = h
        spec = C.c_int()
        check(_lib.lib().fb_model_info(self._h, None, None, None, C.byref(spec)))
        self.specialised = bool(spec.value)

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            try:
                _lib.lib().fb_model_destroy(h)
            except Exception:
                pass
            self._h = None

    # -- helpers ------------------------------------------------------------------------------
    def split(self, flat):
        """[..., N] per-neuron array -> list of per-layer arrays."""
        return [flat[..., self.boff[l]:self.boff[l + 1]] for l in range(self.L)]

    def tolerance(self, lo, hi):
        return sliver_tolerance(self.w, self.b, lo, hi)

    def _dev(self, a, dtype):
        if a is None:
            return None
        if isinstance(a, torch.Tensor):
            if a.device != self.device or a.dtype != dtype or not a.is_contiguous():
                a = a.to(device=self.device, dtype=dtype).contiguous()
            return a
        return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype).to(self.device)

    def _alloc_outputs(self, n, want):
        t = {}
        for k in want:
            if k in MASK_OUTPUTS:
                t[k] = torch.empty((n, self.NW), dtype=torch.int32, device=self.device)
            elif k == "verdict":
                t[k] = torch.empty((n,), dtype=torch.uint8, device=self.device)
            elif k == "counts":
                t[k] = torch.empty((n, self.N), dtype=torch.int16, device=self.device)
            elif k in BOUND_OUTPUTS:
                t[k] = torch.empty((n, self.N), dtype=torch.float64, device=self.device)
            elif k == "record":
                t[k] = torch.empty((n, 2 * self.NW + 2), dtype=torch.int32, device=self.device)
            else:
                raise ValueError(f"unknown output {k!r}")
        po = PruneOut()
        for k in ALL_OUTPUTS:
            setattr(po, k, t[k].data_ptr() if k in t else None)
        return t, po

    @staticmethod
    def _fix_views(t):
        # masks / counts are unsigned on the device; torch lacks uint32/uint16 arithmetic, keep views handy
        return t

    # -- the stage ------------------------------------------------------------------------------
    def profile(self, boxes, S=1000, samples=None, seed=0, first_index=0, out=None):
        """uint32-valued int32 tensor [n, N]: per neuron, the number of samples with a non-zero activation
        (written into `out` when given: contiguous int32 [n, N] on this device)."""
        boxes = self._dev(boxes, torch.int32); n = boxes.shape[0]
        samples = self._dev(samples, torch.int32)
        if samples is not None:
            S = samples.shape[1]
        counts = torch.empty((n, self.N), dtype=torch.int32, device=self.device) if out is None else out
        if out is not None and (out.dtype != torch.int32 or tuple(out.shape) != (n, self.N) or not out.is_contiguous()
                                or out.device != boxes.device):
            raise ValueError("profile: out must be a contiguous int32 [n, N] tensor on the model's device")
        check(_lib.lib().fb_profile(self._h, n, _ptr(boxes), _ptr(samples), int(S), C.c_uint64(seed), int(first_index),
                                    _ptr(counts), _stream()), "fb_profile")
        return counts

    def ibp(self, boxes, counts=None, tol=None, want=("b_dead", "verdict") + BOUND_OUTPUTS):
        boxes = self._dev(boxes, torch.int32); n = boxes.shape[0]
        counts = self._dev(counts, torch.int32); tol = self._dev(tol, torch.float64)
        t, po = self._alloc_outputs(n, want)
        check(_lib.lib().fb_ibp(self._h, n, _ptr(boxes), _ptr(counts), _ptr(tol), C.byref(po), _stream()), "fb_ibp")
        return PruneResult(self, n, None, t)

    def prune(self, boxes, S=1000, samples=None, seed=0, first_index=0, tol=None, want=DEFAULT_OUTPUTS):
        """The whole pre-solver stage on device-resident boxes; outputs stay on the device."""
        boxes = self._dev(boxes, torch.int32); n = boxes.shape[0]
        samples = self._dev(samples, torch.int32)
        if samples is not None:
            S = samples.shape[1]
        tol = self._dev(tol, torch.float64)
        t, po = self._alloc_outputs(n, want)
        check(_lib.lib().fb_prune(self._h, n, _ptr(boxes), _ptr(samples), int(S), C.c_uint64(seed), int(first_index),
                                  _ptr(tol), C.byref(po), _stream()), "fb_prune")
        return PruneResult(self, n, S, t)

    def prune_host(self, boxes, S=1000, samples=None, seed=0, first_index=0, tol=None, want=DEFAULT_OUTPUTS, out=None):
        """Same through HOST buffers (numpy / pinned torch tensors in, pinned torch tensors out)."""
        hb = boxes if isinstance(boxes, torch.Tensor) else torch.as_tensor(np.ascontiguousarray(boxes, dtype=np.int32))
        n = hb.shape[0]
        hs = None
        if samples is not None:
            hs = samples if isinstance(samples, torch.Tensor) else torch.as_tensor(np.ascontiguousarray(samples, dtype=np.int32))
            S = hs.shape[1]
        ht = None if tol is None else torch.as_tensor(np.ascontiguousarray(tol, dtype=np.float64))
        if out is None:
            out = self.alloc_host_outputs(n, want)
        po = PruneOut()
        for k in ALL_OUTPUTS:
            setattr(po, k, out[k].data_ptr() if k in out else None)
        check(_lib.lib().fb_prune_host(self._h, n, _ptr(hb), _ptr(hs), int(S), C.c_uint64(seed), int(first_index),
                                       _ptr(ht), C.byref(po), _stream()), "fb_prune_host")
        return out

    def alloc_host_outputs(self, n, want=DEFAULT_OUTPUTS):
        shapes = {}
        for k in want:
            if k in MASK_OUTPUTS:
                shapes[k] = ((n, self.NW), torch.int32)
            elif k == "verdict":
                shapes[k] = ((n,), torch.uint8)
            elif k == "counts":
                shapes[k] = ((n, self.N), torch.int16)
            elif k == "record":
                shapes[k] = ((n, 2 * self.NW + 2), torch.int32)
            else:
                shapes[k] = ((n, self.N), torch.float64)
        return {k: torch.empty(s, dtype=dt).pin_memory() for k, (s, dt) in shapes.items()}

    # -- logits ---------------------------------------------------------------------------------
    def forward(self, X, dead_bits=None):
        """Logits of `net` for rows X; float64 unless X is float32 (then float32, like numpy)."""
        if isinstance(X, torch.Tensor):
            f32 = X.dtype == torch.float32
        else:
            X = np.asarray(X); f32 = X.dtype == np.float32
        dt = torch.float32 if f32 else torch.float64
        X = self._dev(X, dt).reshape(-1, self.d)
        rows = X.shape[0]
        dead = None
        if dead_bits is not None:
            dead = self._dev(np.asarray(dead_bits, dtype=np.uint32).view(np.int32).reshape(self.NW), torch.int32)
        out = torch.empty((rows,), dtype=dt, device=self.device)
        fn = _lib.lib().fb_forward_f32 if f32 else _lib.lib().fb_forward_f64
        check(fn(self._h, rows, _ptr(X), _ptr(dead), _ptr(out), _stream()), "fb_forward")
        return out

    def heuristic(self, ws_lb, ws_ub, cand, s_cand, deads, pct):
        """(new_deads, merged_deads) int32 mask tensors [n, NW]; inputs [n, N] float64 bounds and [n, NW] masks."""
        ws_lb = self._dev(ws_lb, torch.float64).reshape(-1, self.N); ws_ub = self._dev(ws_ub, torch.float64).reshape(-1, self.N)
        n = ws_lb.shape[0]
        masks = [self._dev(np.asarray(x, dtype=np.uint32).view(np.int32) if not isinstance(x, torch.Tensor) else x,
                           torch.int32).reshape(n, self.NW) for x in (cand, s_cand, deads)]
        new = torch.empty((n, self.NW), dtype=torch.int32, device=self.device)
        merged = torch.empty((n, self.NW), dtype=torch.int32, device=self.device)
        check(_lib.lib().fb_heuristic(self._h, n, _ptr(ws_lb), _ptr(ws_ub), _ptr(masks[0]), _ptr(masks[1]), _ptr(masks[2]),
                                      C.c_double(float(pct)), _ptr(new), _ptr(merged), _stream()), "fb_heuristic")
        return new, merged

    def pair_screen(self, X, pa_cols, alts):
        """(orig_pred uint8 [m], first_flip int32 [m]) for float32 rows X and alternatives [k, n_pa]."""
        X = self._dev(X, torch.float32).reshape(-1, self.d); m = X.shape[0]
        pa = np.ascontiguousarray(pa_cols, dtype=np.int32)
        alts = self._dev(np.asarray(alts, dtype=np.float32).reshape(-1, len(pa)), torch.float32)
        orig = torch.empty((m,), dtype=torch.uint8, device=self.device)
        flip = torch.empty((m,), dtype=torch.int32, device=self.device)
        check(_lib.lib().fb_pair_screen(self._h, m, _ptr(X), len(pa), pa.ctypes.data_as(C.c_void_p), alts.shape[0],
                                        _ptr(alts), _ptr(orig), _ptr(flip), _stream()), "fb_pair_screen")
        return orig, flip

    def certified(self, boxes, ra_col=None, ra_threshold=0):
        """uint8 [n]: 1 where the bounds kernel certifies the logit one-signed over the box -- over the box widened by
        ra_threshold on both sides of column ra_col for a relaxed query -- i.e. where partition_screen need not search."""
        boxes = self._dev(boxes, torch.int32)
        if ra_col is not None and ra_col >= 0 and ra_threshold:
            boxes = boxes.clone()
            boxes[:, ra_col, 0] -= int(ra_threshold); boxes[:, ra_col, 1] += int(ra_threshold)
        v = self.ibp(boxes, want=("verdict",)).t["verdict"]
        return (v & VERDICT_CERTIFIED).to(torch.uint8)

    def partition_screen(self, boxes, pa_col, S=1000, samples=None, seed=0, first_index=0, verdict=None, ra_col=None,
                         ra_threshold=0, record=None, skip=None):
        """Sampled counterexample screen.  Returns [ce_sample, pa_pos, pa_neg] (int32 [n] each), plus [ra_dpos, ra_dneg]
        when a relaxed column is given (relaxed/AC/Verify-AC.py:48-51); `record` = the packed records of prune();
        `skip` = uint8 [n], non-zero where the box is not to be searched (see certified())."""
        boxes = self._dev(boxes, torch.int32); n = boxes.shape[0]
        skip = self._dev(skip, torch.uint8)
        if skip is not None and tuple(skip.shape) != (n,):
            raise ValueError("partition_screen: skip must have one byte per box")
        samples = self._dev(samples, torch.int32)
        if samples is not None:
            S = samples.shape[1]
        relaxed = ra_col is not None and ra_col >= 0
        outs = [torch.empty((n,), dtype=torch.int32, device=self.device) for _ in range(5 if relaxed else 3)]
        check(_lib.lib().fb_partition_screen_ex(self._h, n, _ptr(boxes), _ptr(samples), int(S), C.c_uint64(seed),
                                                int(first_index), int(pa_col), int(ra_col) if relaxed else -1, int(ra_threshold),
                                                _ptr(outs[0]), _ptr(outs[1]), _ptr(outs[2]), _ptr(outs[3]) if relaxed else None,
                                                _ptr(outs[4]) if relaxed else None, _ptr(verdict), _ptr(record), _ptr(skip),
                                                _stream()),
              "fb_partition_screen")
        return outs

    def nearzero(self, boxes, S=1000, samples=None, seed=0, first_index=0):
        """(count, smallest ratio) of |pre-activation| / (sum |w x| + |b|) over every sample and neuron of the batch
        (fb_profile_nearzero): how far the activation signs are from depending on the summation order."""
        boxes = self._dev(boxes, torch.int32); n = boxes.shape[0]
        samples = self._dev(samples, torch.int32)
        if samples is not None:
            S = samples.shape[1]
        cnt = torch.zeros((1,), dtype=torch.int64, device=self.device)
        best = torch.ones((1,), dtype=torch.float64, device=self.device)
        check(_lib.lib().fb_profile_nearzero(self._h, n, _ptr(boxes), _ptr(samples), int(S), C.c_uint64(seed), int(first_index),
                                             _ptr(cnt), _ptr(best), _stream()), "fb_profile_nearzero")
        return int(cnt.item()), float(best.item())


STATUS = {None: 0, "none": 0, "sat": 1, "unsat": 2, "unknown": 3}


def hybrid_predict(grid, original, fairer, X, status):
    """find_partition_result_for_point + hybrid_predict (src/AC/Verify-AC-experiment-new2.py:587-638) for all rows of X at
    once: (float32 logits of the model each row's partition verdict selects, route uint8 [rows], counters int32 [4]).
    status: uint8 [grid size] of STATUS codes (0 = partition not evaluated)."""
    X = np.ascontiguousarray(X)
    x64 = torch.as_tensor(X.astype(np.float64)).to(grid.device); x32 = torch.as_tensor(X.astype(np.float32)).to(grid.device)
    st = status if isinstance(status, torch.Tensor) else torch.as_tensor(np.ascontiguousarray(status, dtype=np.uint8)).to(grid.device)
    rows = x64.shape[0]
    logit = torch.empty((rows,), dtype=torch.float32, device=grid.device)
    route = torch.empty((rows,), dtype=torch.uint8, device=grid.device)
    counters = torch.zeros((4,), dtype=torch.int32, device=grid.device)
    check(_lib.lib().fb_hybrid_predict(grid._h, original._h, fairer._h, rows, _ptr(x64), _ptr(x32), _ptr(st), _ptr(logit),
                                       _ptr(route), _ptr(counters), _stream()), "fb_hybrid_predict")
    return logit, route, counters


class Grid:
    """The partition grid of `partition` + `partitioned_ranges`, addressed by canonical index."""

    def __init__(self, columns, range_dict, T):
        _require_cuda()
        self.columns = list(columns)
        self.keys = list(range_dict.keys())
        if sorted(self.keys) != sorted(self.columns):
            raise ValueError("range_dict keys and columns differ")
        self.T = int(T)
        self.lo = np.array([range_dict[c][0] for c in self.columns], dtype=np.int64)
        self.hi = np.array([range_dict[c][1] for c in self.columns], dtype=np.int64)
        self.key_order = np.array([self.columns.index(k) for k in self.keys], dtype=np.int32)
        torch.cuda.current_stream()
        self.device = torch.device("cuda", torch.cuda.current_device())
        h = C.c_void_p()
        check(_lib.lib().fb_grid_create(len(self.columns), self.lo.ctypes.data_as(C.c_void_p),
                                        self.hi.ctypes.data_as(C.c_void_p), self.T,
                                        self.key_order.ctypes.data_as(C.c_void_p), C.byref(h)), "fb_grid_create")
        self._h = h
        self.size = int(_lib.lib().fb_grid_size(h))

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            try:
                _lib.lib().fb_grid_destroy(h)
            except Exception:
                pass
            self._h = None

    def __len__(self):
        return self.size

    def boxes(self, first=0, count=None, out=None):
        """int32 device tensor [count, d, 2] for canonical indices first .. first+count-1."""
        if count is None:
            count = self.size - first
        if out is None:
            out = torch.empty((count, len(self.columns), 2), dtype=torch.int32, device=self.device)
        check(_lib.lib().fb_grid_boxes(self._h, int(first), int(count), _ptr(out), _stream()), "fb_grid_boxes")
        return out


def _grid_locate(self, X, want_index=True, want_density=False):
    """(index int64 [m] or None, density int32 [size] or None) for rows X [m, d] (column order)."""
    X = torch.as_tensor(np.ascontiguousarray(X, dtype=np.float64)).to(self.device) if not isinstance(X, torch.Tensor) \
        else X.to(device=self.device, dtype=torch.float64).contiguous()
    m = X.shape[0]
    index = torch.empty((m,), dtype=torch.int64, device=self.device) if want_index else None
    density = torch.zeros((self.size,), dtype=torch.int32, device=self.device) if want_density else None
    check(_lib.lib().fb_grid_locate(self._h, m, _ptr(X), _ptr(index), _ptr(density), _stream()), "fb_grid_locate")
    return index, density


Grid.locate = _grid_locate


def fp64_peak(iters=4096):
    """Measured DFMA rate of the current device in TFLOP/s (FMA = 2 flops) and the kernel time in ms."""
    _require_cuda()
    tf, ms = C.c_double(), C.c_double()
    check(_lib.lib().fb_fp64_peak(int(iters), C.byref(tf), C.byref(ms)), "fb_fp64_peak")
    return tf.value, ms.value
