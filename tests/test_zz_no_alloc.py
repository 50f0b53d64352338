"""GPU: fb_prune_noalloc, the prune entry point for embedders that may not allocate device memory.

Some hosts of this library cannot let it call cudaMalloc: a service loop that has sworn off the device
allocator after start-up because an allocation can stall every other stream, a process whose HBM is already
fully handed out, or a framework that counts any allocation it did not make as a leak.  For them the
[n][N] count scratch of fb_prune is a fixed 16 MiB array inside the library.  This file checks, through the
raw C ABI, that the variant returns exactly what fb_prune returns -- on two calls in a row, the second one
finding the first one's stale counters in the shared array -- and that batches over the built-in limit are
refused with a clear message instead of overrunning the array.

The file name sorts last on purpose: a `-x` run has already exercised every other test when it gets here.
It must pass both on its own in a fresh process and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

from tests.common import Golden  # noqa: E402

G = Golden()
NOALLOC_MAX_WORDS = 4 * 1024 * 1024


def _call(fn_name, m, boxes, S, seed, first, tol, want):
    """Run fb_prune or fb_prune_noalloc through ctypes on device tensors; return host copies of the outputs."""
    from fairify_b200 import _lib, engine
    outs, po = m._alloc_outputs(boxes.shape[0], want)
    for t in outs.values():
        # poison: every element must be written by the call
        t.fill_(float("nan") if t.dtype == torch.float64 else (255 if t.dtype == torch.uint8 else -1))
    fn = getattr(_lib.lib(), fn_name)
    rc = fn(m._h, boxes.shape[0], C.c_void_p(boxes.data_ptr()), None, S, C.c_uint64(seed), first,
            C.c_void_p(tol.data_ptr()), C.byref(po), C.c_void_p(torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    return rc, {k: v.cpu().numpy() for k, v in outs.items()}


def _tol(m, name):
    dom = G.domain(name)
    lo = [dom.ranges[c][0] for c in dom.columns]; hi = [dom.ranges[c][1] for c in dom.columns]
    return torch.as_tensor(m.tolerance(lo, hi)).cuda()


def test_noalloc_equals_fb_prune_on_consecutive_calls():
    from fairify_b200 import engine
    dom = G.domain("AC-1")
    m = engine.Model(*G.wb("AC-1"))
    grid = engine.Grid(dom.columns, dom.ranges, 10)
    tol = _tol(m, "AC-1")
    want = engine.ALL_OUTPUTS
    # call 1: one warp per partition (counters are stored); call 2: three partitions, so the samples are
    # split over many warps and the counters are accumulated atomically on top of call 1's leftovers;
    # call 3: a middle-sized batch again, over the array both earlier calls dirtied
    plan = [(0, 16000, 200, 17), (5, 3, 1000, 18), (9000, 4097, 96, 19)]
    for first, n, S, seed in plan:
        boxes = grid.boxes(first, n)
        rc_a, a = _call("fb_prune", m, boxes, S, seed, first, tol, want)
        rc_b, b = _call("fb_prune_noalloc", m, boxes, S, seed, first, tol, want)
        assert rc_a == 0 and rc_b == 0
        for k in want:
            assert np.array_equal(a[k], b[k], equal_nan=False), f"call (first={first}, n={n}): output {k} differs"


def test_noalloc_runtime_shape_model_and_limit():
    from fairify_b200 import _lib, engine
    name = "BM-4"                                  # served by the runtime-shape kernels; N = 301
    dom = G.domain(name)
    m = engine.Model(*G.wb(name))
    assert not m.specialised
    grid = engine.Grid(dom.columns, dom.ranges, dom.threshold)
    tol = _tol(m, name)
    want = ("cand", "b_dead", "dead", "sliver", "verdict", "counts", "ws_ub", "ub_ru")
    for first, n, S, seed in [(0, 64, 64, 3), (100, 7, 300, 4)]:
        boxes = grid.boxes(first, n)
        rc_a, a = _call("fb_prune", m, boxes, S, seed, first, tol, want)
        rc_b, b = _call("fb_prune_noalloc", m, boxes, S, seed, first, tol, want)
        assert rc_a == 0 and rc_b == 0
        for k in want:
            assert np.array_equal(a[k], b[k]), k
    # one partition too many for the built-in array: refused, nothing launched, outputs untouched
    n_over = NOALLOC_MAX_WORDS // m.N + 1
    boxes = torch.zeros((n_over, m.d, 2), dtype=torch.int32, device="cuda")
    rc, out = _call("fb_prune_noalloc", m, boxes, 8, 1, 0, tol, ("verdict",))
    assert rc != 0
    msg = _lib.lib().fb_last_error().decode()
    assert "exceed the built-in scratch" in msg and str(NOALLOC_MAX_WORDS) in msg
    assert np.all(out["verdict"] == 255)
    # exactly at the limit is accepted
    n_ok = NOALLOC_MAX_WORDS // m.N
    rc, out = _call("fb_prune_noalloc", m, boxes[:n_ok], 8, 1, 0, tol, ("verdict",))
    assert rc == 0 and not np.any(out["verdict"] == 255)


def test_noalloc_call_allocates_nothing():
    """After a warm-up call (which resolves the scratch array's address), a call must leave the device's free memory
    and the default stream-ordered pool exactly as they were -- fb_prune, by contrast, grows this library's pool."""
    from fairify_b200 import _lib, engine
    cudart = C.CDLL("libcudart.so.12") if hasattr(C, "CDLL") else None
    dom = G.domain("AC-1")
    m = engine.Model(*G.wb("AC-1"))
    grid = engine.Grid(dom.columns, dom.ranges, 10)
    tol = _tol(m, "AC-1")
    boxes = grid.boxes(0, 12000)
    want = ("cand", "b_dead", "dead", "verdict")
    outs, po = m._alloc_outputs(boxes.shape[0], want)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def call():
        rc = _lib.lib().fb_prune_noalloc(m._h, boxes.shape[0], C.c_void_p(boxes.data_ptr()), None, 64, C.c_uint64(5), 0,
                                         C.c_void_p(tol.data_ptr()), C.byref(po), st)
        torch.cuda.synchronize()
        return rc

    def pool_bytes():
        pool = C.c_void_p()
        assert cudart.cudaDeviceGetDefaultMemPool(C.byref(pool), torch.cuda.current_device()) == 0
        used, reserved = C.c_uint64(), C.c_uint64()
        # cudaMemPoolAttrReservedMemCurrent = 5, cudaMemPoolAttrUsedMemCurrent = 7
        assert cudart.cudaMemPoolGetAttribute(pool, 5, C.byref(reserved)) == 0
        assert cudart.cudaMemPoolGetAttribute(pool, 7, C.byref(used)) == 0
        return used.value, reserved.value

    assert call() == 0                                  # warm-up: address lookup, module load
    free0 = torch.cuda.mem_get_info()[0]; pool0 = pool_bytes()
    for _ in range(3):
        assert call() == 0
    assert torch.cuda.mem_get_info()[0] == free0 and pool_bytes() == pool0
    # a model on another device than the current one is refused (the scratch address belongs to the current device)
    if torch.cuda.device_count() > 1:
        with torch.cuda.device(1):
            assert call() != 0 and b"device" in _lib.lib().fb_last_error()
