"""GPU: the CUDA path (through the C ABI) against the oracle and the reference's golden vectors.

Bar: bit-exact for masks, counts, verdicts, grid boxes, samples and the round-to-nearest bounds;
the outward-rounded bounds must contain the reference bounds and stay within 1e-9 relative of them."""
import hashlib
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

from oracle import oracle as O  # noqa: E402
from tests.common import Golden  # noqa: E402

G = Golden()
MODELS = G.models


@pytest.fixture(scope="module")
def eng():
    from fairify_b200 import engine
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return engine


def _tol(eng, name):
    dom = G.domain(name)
    lo = np.array([dom.ranges[c][0] for c in dom.columns]); hi = np.array([dom.ranges[c][1] for c in dom.columns])
    w, b = G.wb(name)
    return eng.sliver_tolerance(w, b, lo, hi)


def test_library_loaded_and_native(eng):
    from fairify_b200 import _lib
    assert _lib.lib().fb_version() >= 100
    m = eng.Model(*G.wb("AC-1"))
    assert m.specialised and m.N == 25 and m.P == 344


@pytest.mark.parametrize("tag", ["AC_T10", "GC_T100", "BM_T100", "GC_T10", "CP_T5", "AC_T6_age30_35",
                                 "GC_T10_purpose7", "BM_T10_small"])
def test_grid_boxes_match_reference(eng, tag):
    from fairify_b200.domains import DOMAINS
    g = G.grids
    dom = DOMAINS[tag.split("_")[0]]
    rd = dom.copy_ranges()
    lo, hi = g[f"{tag}/lo"], g[f"{tag}/hi"]
    for c, a, b in zip(dom.columns, lo, hi):
        rd[c] = [int(a), int(b)]
    grid = eng.Grid(dom.columns, rd, int(g[f"{tag}/T"]))
    assert len(grid) == int(g[f"{tag}/n"])
    boxes = grid.boxes().cpu().numpy()
    assert hashlib.sha256(boxes.tobytes()).digest() == g[f"{tag}/sha256"].tobytes()
    # ragged windows: a slice in the middle and the empty slice
    mid = grid.boxes(len(grid) // 3, min(5, len(grid) - len(grid) // 3)).cpu().numpy()
    assert np.array_equal(mid, boxes[len(grid) // 3: len(grid) // 3 + mid.shape[0]])
    assert grid.boxes(0, 0).shape[0] == 0
    with pytest.raises(Exception):
        grid.boxes(len(grid) - 1, 2)


@pytest.mark.parametrize("name", MODELS)
def test_prune_matches_reference_goldens(eng, name):
    """sound_prune* outputs on the reference's own samples (injected), all golden cases of the model."""
    w, b = G.wb(name)
    m = eng.Model(w, b)
    tol = _tol(eng, name)
    by_S = {}
    for c in G.cases(name):
        by_S.setdefault(int(c["S"]), []).append(c)
    for S, cs in by_S.items():
        boxes = np.stack([c["box"] for c in cs]); samples = np.stack([c["samples"] for c in cs])
        r = m.prune(boxes, samples=samples, tol=tol, want=eng.ALL_OUTPUTS)
        cand, b_dead, dead, sliver = r.bits("cand"), r.bits("b_dead"), r.bits("dead"), r.bits("sliver")
        s_dead, s_cand = r.bits("s_dead"), r.bits("s_cand")
        counts = r.numpy("counts").view(np.uint16)
        # where the fixture has no Z3 run, the oracle's closed form (pinned by tests/test_oracle_golden.py) stands in
        net = O.Net(w, b)
        om = O.masks(net, np.stack([O.profile(net, c["samples"][None])[0] for c in cs]),
                     np.stack([c["ws_ub"] for c in cs]), np.stack([c["pl_lb"] for c in cs]),
                     np.stack([c["pl_ub"] for c in cs]), tol)
        for k, c in enumerate(cs):
            assert np.array_equal(cand[k], c["cand"]), c["tag"]
            assert np.array_equal(counts[k] / S, c["pos_prob"]), c["tag"]
            assert np.array_equal(b_dead[k], c["b_dead"]), c["tag"]
            assert not sliver[k].any()
            if "dead" in c:
                assert np.array_equal(dead[k], c["dead"]), c["tag"]
                assert np.array_equal(s_dead[k], c["s_dead"]) and np.array_equal(s_cand[k], c["s_cand"]), c["tag"]
            assert np.array_equal(dead[k], om["dead"][k]) and np.array_equal(s_dead[k], om["s_dead"][k]), c["tag"]
            assert np.array_equal(s_cand[k], om["s_cand"][k]), c["tag"]
            for key in ("ws_lb", "ws_ub", "pl_lb", "pl_ub"):
                assert np.array_equal(r.numpy(key)[k], c[key]), f"{c['tag']} {key}"
            lb_rd, ub_ru = r.numpy("lb_rd")[k], r.numpy("ub_ru")[k]
            assert np.all(lb_rd <= c["ws_lb"]) and np.all(ub_ru >= c["ws_ub"]), c["tag"]
            scale = np.maximum(1.0, np.maximum(np.abs(c["ws_lb"]), np.abs(c["ws_ub"])))
            assert np.all(c["ws_lb"] - lb_rd <= 1e-9 * scale) and np.all(ub_ru - c["ws_ub"] <= 1e-9 * scale)


@pytest.mark.parametrize("name", G.singular_models)
def test_singular_verification_matches_reference_z3(eng, name):
    """s_dead / s_candidates / dead_nodes of the full sound_prune* INCLUDING the Z3 singular_verification
    (utils/prune.py:276-644), on the fixtures of every model with an equal-width hidden layer: the device's closed
    form of the query must reproduce what the unmodified reference + Z3 returned (neurons declared dead although their
    interval bound is positive included), and the packed record must carry the same masks."""
    w, b = G.wb(name)
    m = eng.Model(w, b)
    tol = _tol(eng, name)
    cs = list(G.singular_cases(name))
    boxes = np.stack([c["box"] for c in cs]); samples = np.stack([c["samples"] for c in cs])
    r = m.prune(boxes, samples=samples, tol=tol, want=eng.ALL_OUTPUTS)
    rec = r.numpy("record").view(np.uint32)
    for key in ("cand", "b_dead", "s_dead", "s_cand", "dead"):
        got = r.bits(key)
        for k, c in enumerate(cs):
            assert np.array_equal(got[k], c[key]), f"{c['tag']} {key}"
    assert not r.bits("sliver").any()
    assert np.array_equal(r.numpy("ws_ub"), np.stack([c["ws_ub"] for c in cs]))
    assert np.array_equal(rec[:, 0], r.numpy("verdict").astype(np.uint32))
    assert np.array_equal(rec[:, 1:1 + m.NW], r.numpy("dead").view(np.uint32))
    assert np.array_equal(rec[:, 1 + m.NW:1 + 2 * m.NW], r.numpy("cand").view(np.uint32))
    assert np.all(rec[:, -1] == 0xffffffff)
    hits = sum(int(c["s_dead"].sum()) for c in cs)
    if name in ("AC-4", "AC-8", "AC-10", "AC-12", "BM-6", "BM-7"):
        assert hits > 0 and r.bits("s_dead").sum() == hits


@pytest.mark.parametrize("name", ["AC-1", "AC-11", "GC-4", "GC-1"])
def test_bounds_on_boxes_beyond_the_exact_product_range(eng, name):
    """The looped bounds kernel fuses product and sum in layer 0 only while |x| < 2^29 keeps the products exact; boxes
    beyond that take its unfused path.  Both must give the oracle's bounds bit for bit, mixed in one batch."""
    w, b = G.wb(name)
    m = eng.Model(w, b); net = O.Net(w, b)
    dom = G.domain(name)
    base = np.array([[dom.ranges[c][0], dom.ranges[c][1]] for c in dom.columns], dtype=np.int32)
    boxes = np.repeat(base[None], 6, axis=0).copy()
    boxes[1, 0] = [-(2 ** 29) - 5, 2 ** 29 + 7]
    boxes[2, 2] = [2 ** 30, 2 ** 31 - 1]
    boxes[3, 1] = [-(2 ** 31), -(2 ** 30)]
    boxes[4, 0] = [2 ** 29 - 1, 2 ** 29 - 1]
    boxes[5, 0] = [2 ** 29, 2 ** 29]
    r = m.ibp(boxes, want=eng.BOUND_OUTPUTS + ("b_dead", "verdict"))
    o_rn = O.ibp(net, boxes); o_dir = O.ibp_directed(net, boxes)
    for key, ref in zip(("ws_lb", "ws_ub", "pl_lb", "pl_ub"), o_rn):
        assert np.array_equal(r.numpy(key), ref), key
    assert np.array_equal(r.numpy("lb_rd"), o_dir[0]) and np.array_equal(r.numpy("ub_ru"), o_dir[1])


def test_sliver_flag_set_by_a_constructed_case(eng):
    """A network built so that the singular-verification quantity of a hidden neuron lands inside the tolerance: the
    device must flag it (sliver bit + FB_VERDICT_SLIVER), exactly like the oracle, on layer 0 (quantity = ws_ub) and
    on a square layer 1 (quantity = q)."""
    # layer 0 neuron 0 over the box [0,4]x[0,4]: 2^-52 * 4 + 1 * 4 - 4 = 2^-50, far inside the tolerance of sums of
    # magnitude 8; the other neurons are plainly positive (with a forced tolerance of 1e3 they are all flagged, which
    # exercises the |q| <= tol test of the square layer 1)
    w = [np.array([[2.0 ** -52, 1.0], [1.0, 1.0]], np.float32), np.array([[0.5, 0.25], [-0.5, 0.5]], np.float32),
         np.array([[1.0], [1.0]], np.float32)]
    b = [np.array([-4.0, 0.0], np.float32), np.array([0.125, -1.0], np.float32), np.array([0.25], np.float32)]
    m = eng.Model(w, b); net = O.Net(w, b)
    box = np.array([[[0, 4], [0, 4]]], dtype=np.int32)
    tol = eng.sliver_tolerance(w, b, box[0, :, 0], box[0, :, 1])
    assert np.array_equal(tol, O.sliver_tolerance(net, box[0, :, 0], box[0, :, 1]))
    for force_tol in (tol, np.full_like(tol, 1e3)):          # the honest tolerance, then one that flags every candidate
        r = m.ibp(box, tol=force_tol, want=eng.ALL_OUTPUTS)
        _, ws_ub, pl_lb, pl_ub = O.ibp(net, box)
        om = O.masks(net, None, ws_ub, pl_lb, pl_ub, force_tol)
        for key in ("cand", "b_dead", "s_dead", "s_cand", "dead", "sliver"):
            assert np.array_equal(r.bits(key), om[key]), key
        assert bool(r.numpy("verdict")[0] & eng.VERDICT_SLIVER) == bool(om["sliver"].any())
    assert om["sliver"].any()
    r = m.ibp(box, tol=tol, want=("sliver",))
    assert r.bits("sliver")[0][0] == 1 and r.bits("sliver").sum() == 1      # layer 0 neuron 0: 0 < ws_ub = 2^-50 <= tol


@pytest.mark.parametrize("name", MODELS)
def test_prune_matches_oracle_philox(eng, name):
    """device-drawn samples: counts, masks, verdict and both bound chains against the oracle, bit for bit."""
    from fairify_b200.domains import STRESS_THRESHOLD
    w, b = G.wb(name)
    m = eng.Model(w, b); net = O.Net(w, b)
    dom = G.domain(name)
    T = STRESS_THRESHOLD.get(dom.name, dom.threshold)
    grid = eng.Grid(dom.columns, dom.ranges, T)
    big = m.P > 4000
    n = 24 if big else 96
    first = max(0, len(grid) // 2 - 7)
    n = min(n, len(grid) - first)
    S = 100 if big else 333                      # ragged: not a multiple of 32
    boxes = grid.boxes(first, n)
    hb = boxes.cpu().numpy()
    tol = _tol(eng, name)
    r = m.prune(boxes, S=S, seed=99, first_index=first, tol=tol, want=eng.ALL_OUTPUTS)
    want = O.prune_batch(net, hb, 99, first, S, tol)
    assert np.array_equal(r.numpy("counts").view(np.uint16), want["counts"].astype(np.uint16))
    assert np.array_equal(r.bits("cand"), want["cand"])
    assert np.array_equal(r.bits("b_dead"), want["b_dead"])
    assert np.array_equal(r.bits("dead"), want["dead"])
    assert np.array_equal(r.bits("s_dead"), want["s_dead"])
    assert np.array_equal(r.bits("sliver"), want["sliver"])
    assert np.array_equal(r.numpy("verdict"), want["verdict"])
    o_rn = O.ibp(net, hb); o_dir = O.ibp_directed(net, hb)
    for key, ref in zip(("ws_lb", "ws_ub", "pl_lb", "pl_ub"), o_rn):
        assert np.array_equal(r.numpy(key), ref), key
    assert np.array_equal(r.numpy("lb_rd"), o_dir[0]) and np.array_equal(r.numpy("ub_ru"), o_dir[1])


@pytest.mark.parametrize("name", ["AC-1", "GC-1", "BM-4"])
def test_profile_few_partitions_uses_sample_chunks(eng, name):
    """with very few partitions the samples are split over several warps and merged atomically."""
    w, b = G.wb(name)
    m = eng.Model(w, b); net = O.Net(w, b)
    dom = G.domain(name)
    grid = eng.Grid(dom.columns, dom.ranges, dom.threshold)
    boxes = grid.boxes(0, 3)
    counts = m.profile(boxes, S=1000, seed=5, first_index=0).cpu().numpy()
    samples = O.philox_samples(5, 0, boxes.cpu().numpy(), 1000)
    assert np.array_equal(counts.astype(np.uint32), O.profile(net, samples))
    # injected copy of the same samples must give the same counts
    counts2 = m.profile(boxes, samples=samples).cpu().numpy()
    assert np.array_equal(counts, counts2)


def _mixed_class_boxes(d, n, seed):
    """Synthetic boxes (not tied to a dataset's domain) whose widest range puts them, irregularly interleaved, in each
    of the three draw classes: 4, 2 and 1 features per Philox word."""
    rng = np.random.RandomState(seed)
    cap = np.array([16, 256, 5000])[rng.randint(0, 3, size=n)]                 # class of each box
    lo = rng.randint(-40, 60, size=(n, d))
    width = 1 + (rng.rand(n, d) * cap[:, None]).astype(np.int64)
    width[np.arange(n), rng.randint(0, d, size=n)] = cap                         # one feature at the class limit
    return np.stack([lo, lo + width - 1], axis=2).astype(np.int32)


@pytest.mark.parametrize("name", ["AC-1", "GC-1", "BM-1", "AC-7", "BM-4"])
def test_profile_mixed_draw_classes(eng, name):
    """one batch holding boxes of all three draw classes: every partition is handled exactly once, by the launch of
    its own class, and matches the oracle's samples (specialised and runtime-shape kernels, partition screen too)."""
    w, b = G.wb(name)
    m = eng.Model(w, b); net = O.Net(w, b)
    dom = G.domain(name)
    n, S = (61 if m.P > 4000 else 301), 70
    hb = _mixed_class_boxes(m.d, n, 17)
    widest = (hb[:, :, 1] - hb[:, :, 0] + 1).max(axis=1)
    assert (widest <= 16).any() and ((widest > 16) & (widest <= 256)).any() and (widest > 256).any()
    boxes = torch.as_tensor(hb).cuda()
    samples = O.philox_samples(21, 1000, hb, S)
    assert np.all(samples >= hb[:, None, :, 0]) and np.all(samples <= hb[:, None, :, 1])
    counts = torch.full((n, m.N), -1, dtype=torch.int32, device="cuda")          # poisoned: every row must be written
    counts = m.profile(boxes, S=S, seed=21, first_index=1000, out=counts).cpu().numpy()
    assert np.array_equal(counts.astype(np.uint32), O.profile(net, samples))
    pa_col = dom.columns.index(dom.pa[0])
    ce = m.partition_screen(boxes, pa_col, S=S, seed=21, first_index=1000)[0].cpu().numpy()
    assert np.array_equal(ce, O.partition_screen(net, samples, hb, pa_col)[0])


@pytest.mark.parametrize("ublock", ["32", "5", None])
def test_profile_block_scheduling_large_batch(eng, ublock, monkeypatch):
    """work handed out in blocks of consecutive partitions (forced to 32 and to an odd size through the test hook, then
    the launcher's own choice), mixed classes, ragged tail: counts against the oracle."""
    if ublock is None:
        monkeypatch.delenv("FAIRIFY_B200_UBLOCK", raising=False)
    else:
        monkeypatch.setenv("FAIRIFY_B200_UBLOCK", ublock)
    w, b = G.wb("AC-1")
    m = eng.Model(w, b); net = O.Net(w, b)
    dom = G.domain("AC-1")
    n, S = 40_003, 40
    hb = _mixed_class_boxes(m.d, n, 23)
    hb[: n // 2, :, 1] = np.minimum(hb[: n // 2, :, 1], hb[: n // 2, :, 0] + 3)   # first half: one class only, second half mixed
    boxes = torch.as_tensor(hb).cuda()
    counts = torch.full((n, m.N), -1, dtype=torch.int32, device="cuda")
    counts = m.profile(boxes, S=S, seed=8, first_index=5, out=counts).cpu().numpy()
    want = O.profile(net, O.philox_samples(8, 5, hb, S))
    assert np.array_equal(counts.astype(np.uint32), want)


def test_subnormal_capable_model_takes_exact_sign_test(eng):
    """The specialised profile kernel reads a pre-activation's sign from its high word, which is only exact when no
    positive subnormal can occur; a 10-layer net of subnormal float32 weights could produce them, so it must be routed
    to the runtime-shape kernels -- and still match the oracle, whose compare is exact."""
    dims = [13, 5, 5, 5, 5, 5, 5, 5, 5, 5, 1]
    rng = np.random.RandomState(3)
    tiny = np.float32(2.0 ** -136)             # the last hidden layer's pre-activations land near 2^-1056
    w = [(rng.randint(1, 4, size=(a, b)).astype(np.float32) * (np.float32(1.0) if l == 0 else tiny))
         for l, (a, b) in enumerate(zip(dims[:-1], dims[1:]))]
    b = [np.zeros(n, dtype=np.float32) for n in dims[1:]]
    m = eng.Model(w, b); net = O.Net(w, b)
    assert not m.specialised
    ok = eng.Model(*G.wb("AC-12"))
    assert ok.specialised                      # the real AC-12 (same shape) keeps the specialised kernels
    dom = G.domain("AC-12")
    boxes = eng.Grid(dom.columns, dom.ranges, dom.threshold).boxes(0, 50)
    counts = m.profile(boxes, S=64, seed=4).cpu().numpy()
    want = O.profile(net, O.philox_samples(4, 0, boxes.cpu().numpy(), 64))
    assert np.array_equal(counts.astype(np.uint32), want)
    assert want[:, 40:45].all()                # the last hidden layer is active with values below 2^-1042


@pytest.mark.parametrize("name,tiled,mma", [
    ("AC-1", "1", "1"), ("AC-12", "1", "1"), ("GC-2", "1", "1"), ("BM-1", "1", "1"), ("GC-5", "1", "1"), ("AC-7", "1", "1"),
    ("GC-3", "1", "1"), ("AC-9", "1", "1"), ("BM-4", "1", "1"), ("AC-4", "1", "1"), ("BM-7", "1", "1"),
    ("AC-1", "1", "0"), ("BM-1", "1", "0"), ("BM-4", "1", "0"), ("AC-5", "1", "0"),
    ("AC-1", "0", "0"), ("BM-1", "0", "0"), ("BM-4", "0", "0"), ("AC-5", "0", "0"), ("AC-10", "0", "0"), ("AC-8", "0", "0"),
    ("BM-6", "0", "0")])
def test_generic_kernels_agree_with_specialised(eng, name, tiled, mma):
    """FAIRIFY_B200_FORCE_GENERIC routes a specialised shape through the runtime-shape kernels (subprocess); the profile
    goes through the DMMA kernel (FAIRIFY_B200_MMA_PROFILE=1), the older tiled kernel (=0 with
    FAIRIFY_B200_TILED_PROFILE=1) or the sample-per-lane kernel (both 0)."""
    import subprocess
    import sys
    code = f"""
import numpy as np, sys
sys.path.insert(0, {os.path.dirname(os.path.dirname(os.path.abspath(__file__)))!r})
from tests.common import Golden
from fairify_b200 import engine
from oracle import oracle as O
G = Golden(); w, b = G.wb({name!r}); m = engine.Model(w, b); net = O.Net(w, b)
dom = G.domain({name!r}); grid = engine.Grid(dom.columns, dom.ranges, dom.threshold)
n = min(40, len(grid)); boxes = grid.boxes(0, n); hb = boxes.cpu().numpy()
lo = np.array([dom.ranges[c][0] for c in dom.columns]); hi = np.array([dom.ranges[c][1] for c in dom.columns])
tol = engine.sliver_tolerance(w, b, lo, hi)
r = m.prune(boxes, S=77, seed=3, tol=tol, want=engine.ALL_OUTPUTS)
want = O.prune_batch(net, hb, 3, 0, 77, tol)
assert np.array_equal(r.numpy('counts').view(np.uint16), want['counts'].astype(np.uint16))
assert np.array_equal(r.bits('b_dead'), want['b_dead']) and np.array_equal(r.numpy('verdict'), want['verdict'])
assert np.array_equal(r.bits('s_dead'), want['s_dead']) and np.array_equal(r.bits('dead'), want['dead'])
if {name!r} in G.singular_models:
    cs = list(G.singular_cases({name!r}))
    rs = m.prune(np.stack([c['box'] for c in cs]), samples=np.stack([c['samples'] for c in cs]), tol=tol, want=engine.ALL_OUTPUTS)
    for key in ('cand', 'b_dead', 's_dead', 's_cand', 'dead'):
        assert np.array_equal(rs.bits(key), np.stack([c[key] for c in cs])), key
    rec = rs.numpy('record').view(np.uint32)
    assert np.array_equal(rec[:, 1:1 + m.NW], rs.numpy('dead').view(np.uint32)) and np.all(rec[:, -1] == 0xffffffff)
    assert np.array_equal(rec[:, 0], rs.numpy('verdict').astype(np.uint32))
assert np.array_equal(r.numpy('ws_ub'), O.ibp(net, hb)[1]) and np.array_equal(r.numpy('ub_ru'), O.ibp_directed(net, hb)[1])
X = O.philox_samples(3, 0, hb, 8).reshape(-1, m.d)
assert np.array_equal(m.forward(X.astype(np.float64)).cpu().numpy(), O.forward_f64(net, X))
assert np.array_equal(m.forward(X.astype(np.float32)).cpu().numpy(), O.forward_f32(net, X))
pa = dom.columns.index(dom.pa[0])
ce = m.partition_screen(boxes, pa, S=77, seed=3)[0].cpu().numpy()
assert np.array_equal(ce, O.partition_screen(net, O.philox_samples(3, 0, hb, 77), hb, pa)[0])
print('generic-ok', m.specialised)
"""
    env = dict(os.environ, FAIRIFY_B200_FORCE_GENERIC="1", FAIRIFY_B200_TILED_PROFILE=tiled, FAIRIFY_B200_MMA_PROFILE=mma)
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "generic-ok" in out.stdout, out.stdout + out.stderr


@pytest.mark.parametrize("name", MODELS)
def test_forward_logits(eng, name):
    """net: float64 on integer rows, float32 on float32 rows; bit-exact vs the oracle, sign-exact vs numpy."""
    w, b = G.wb(name)
    m = eng.Model(w, b); net = O.Net(w, b)
    for c in list(G.cases(name))[:3]:
        rows = c["samples"][:64]
        y64 = m.forward(rows.astype(np.float64)).cpu().numpy()
        y32 = m.forward(rows.astype(np.float32)).cpu().numpy()
        assert np.array_equal(y64, O.forward_f64(net, rows)) and np.array_equal(y32, O.forward_f32(net, rows))
        assert np.array_equal(y64 > 0, c["logit64"] > 0) and np.array_equal(y32 > 0, c["logit32"] > 0)
        assert np.allclose(y64, c["logit64"], rtol=1e-12, atol=1e-12)


@pytest.mark.parametrize("name", ["AC-1", "GC-1", "AC-4", "BM-8"])
def test_forward_with_dead_mask_equals_pruned_network(eng, name):
    """zeroing dead activations == prune_neurons' column/row deletion (utils/prune.py:950-977)."""
    w, b = G.wb(name)
    m = eng.Model(w, b)
    c = next(iter(G.cases(name)))
    dead = c["b_dead"].astype(bool)
    layers = m.split(dead)
    pw, pb = [x.copy() for x in w], [x.copy() for x in b]
    for l in range(m.L):
        keep = ~layers[l]
        pw[l] = pw[l][:, keep]; pb[l] = pb[l][keep]
        if l + 1 < m.L:
            pw[l + 1] = pw[l + 1][keep, :]
    rows = c["samples"][:128].astype(np.float64)
    y_masked = m.forward(rows, dead_bits=eng.pack_bits(c["b_dead"][None])[0]).cpu().numpy()
    y_pruned = eng.Model(pw, pb).forward(rows).cpu().numpy()
    assert np.array_equal(y_masked, y_pruned)
    # sound pruning never changes a prediction on the partition's own samples ('Pruned-acc' == 1.0)
    assert np.array_equal(y_masked > 0, m.forward(rows).cpu().numpy() > 0)


@pytest.mark.parametrize("ci", range(6))
def test_pair_screen_matches_reference_detector(eng, ci):
    """CausalDiscriminationDetector.causal_discrimination (src/AC/metrics.py:101-168): rate, test cases and
    causal pairs reproduced from device predictions + the host replay of the sampler and stopping rule."""
    from fairify_b200.screen import causal_discrimination
    from fairify_b200.domains import domain_for_model
    z = G.causal
    name = str(z[f"case{ci}/model"]); pa = [str(x) for x in z[f"case{ci}/pa"]]
    dom = domain_for_model(name)
    m = eng.Model(*G.wb(name))
    values = {c: [float(v) for v in z[f"case{ci}/values/{c}"]] for c in dom.columns}
    test_cases, rate, pairs = causal_discrimination(m, dom.columns, values, pa, conf=float(z[f"case{ci}/conf"]),
                                                    margin=float(z[f"case{ci}/margin"]))
    assert rate == float(z[f"case{ci}/rate"])
    assert np.array_equal(np.asarray(test_cases, dtype=np.float32), z[f"case{ci}/test_cases"])
    assert np.array_equal(np.asarray(pairs, dtype=np.float32).reshape(-1, 2, dom.d), z[f"case{ci}/pairs"])


@pytest.mark.parametrize("name", ["AC-1", "AC-3", "GC-1", "GC-4", "BM-5", "AC-5", "BM-4", "BM-7", "AC-4", "GC-5"])
def test_partition_screen_matches_oracle(eng, name):
    w, b = G.wb(name)
    m = eng.Model(w, b); net = O.Net(w, b)
    dom = G.domain(name)
    grid = eng.Grid(dom.columns, dom.ranges, dom.threshold)
    n = min(200, len(grid)) if m.P < 4500 else min(60, len(grid))
    boxes = grid.boxes(0, n); hb = boxes.cpu().numpy()
    # the protected attribute, the first column, and the column with the widest range in these boxes (a sweep of up to
    # `threshold` values: wider than one tile of the tiled kernel for the T=100 grids)
    widest = dom.columns[int((hb[:, :, 1] - hb[:, :, 0]).max(axis=0).argmax())]
    for pa_name in dict.fromkeys((dom.pa[0], dom.columns[0], widest)):
        pa_col = dom.columns.index(pa_name)
        S = 150
        verdict = torch.zeros(n, dtype=torch.uint8, device="cuda")
        ce, vp, vn = [t.cpu().numpy() for t in m.partition_screen(boxes, pa_col, S=S, seed=11, verdict=verdict)]
        samples = O.philox_samples(11, 0, hb, S)
        oce, ovp, ovn = O.partition_screen(net, samples, hb, pa_col)
        assert np.array_equal(ce, oce) and np.array_equal(vp, ovp) and np.array_equal(vn, ovn)
        assert np.array_equal(verdict.cpu().numpy() == 2, oce >= 0)
        # injected samples give the same answer
        ce2, vp2, vn2 = [t.cpu().numpy() for t in m.partition_screen(boxes, pa_col, samples=samples)]
        assert np.array_equal(ce2, oce) and np.array_equal(vp2, ovp) and np.array_equal(vn2, ovn)
        # every reported counterexample really flips the float32 class
        for p in np.nonzero(oce >= 0)[0][:20]:
            x = samples[p, oce[p]].astype(np.float32)
            xa, xb = x.copy(), x.copy(); xa[pa_col] = ovp[p]; xb[pa_col] = ovn[p]
            ya, yb = O.forward_f32(net, np.stack([xa, xb]))
            assert ya > 0 and yb < 0


@pytest.mark.parametrize("name", ["AC-1", "AC-3", "AC-5", "AC-4", "AC-10", "AC-7", "AC-2"])
def test_relaxed_partition_screen_matches_oracle(eng, name):
    """relaxed/AC query (PA = race, RA = age, threshold 5; relaxed/AC/Verify-AC.py:48-51): counterexample sample, protected
    values and the difference applied to the relaxed column, bit for bit against the oracle; every reported pair is
    re-checked to flip the float32 class; with no relaxed column the answers are those of the strict screen."""
    w, b = G.wb(name)
    m = eng.Model(w, b); net = O.Net(w, b)
    dom = G.domain(name)
    pa, ra, thr = dom.columns.index("race"), dom.columns.index("age"), 5
    grid = eng.Grid(dom.columns, dom.ranges, 6)                    # stress / relaxed grid of the Adult drivers
    first = len(grid) // 3
    n = 48 if m.P > 4000 else 160
    boxes = grid.boxes(first, n); hb = boxes.cpu().numpy()
    S = 70
    samples = O.philox_samples(11, first, hb, S)
    want = O.partition_screen_ex(net, samples, hb, pa, ra, thr)
    verdict = torch.zeros((n,), dtype=torch.uint8, device="cuda")
    got = m.partition_screen(boxes, pa, S=S, seed=11, first_index=first, ra_col=ra, ra_threshold=thr, verdict=verdict)
    for g_, w_, key in zip(got, want, ("ce_sample", "pa_pos", "pa_neg", "ra_dpos", "ra_dneg")):
        assert np.array_equal(g_.cpu().numpy(), w_), key
    ce = want[0]
    assert np.array_equal((verdict.cpu().numpy() & eng.VERDICT_CE) != 0, ce >= 0)
    for p in np.flatnonzero(ce >= 0)[:40]:
        x = samples[p, ce[p]].astype(np.float32)
        xp = x.copy(); xp[pa] = want[1][p]; xp[ra] += want[3][p]
        xn = x.copy(); xn[pa] = want[2][p]; xn[ra] += want[4][p]
        y = m.forward(np.stack([xp, xn])).cpu().numpy()
        assert y[0] > 0 > y[1]
    # injected samples give the same; no relaxed column = the strict screen
    got2 = m.partition_screen(boxes, pa, samples=samples, ra_col=ra, ra_threshold=thr)
    assert all(np.array_equal(a.cpu().numpy(), b_) for a, b_ in zip(got2, want))
    strict = m.partition_screen(boxes, pa, S=S, seed=11, first_index=first)
    assert np.array_equal(strict[0].cpu().numpy(), O.partition_screen(net, samples, hb, pa)[0])
    assert (ce >= 0).sum() >= (strict[0].cpu().numpy() >= 0).sum()


@pytest.mark.parametrize("name,thr", [("AC-5", 20), ("AC-4", 13), ("AC-1", 20)])
def test_relaxed_screen_with_a_sweep_wider_than_a_tile(eng, name, thr):
    """A relaxed sweep of W * (2 thr + 1) rows per sample that does not fit one tile of the big networks' screen kernel
    (cut into segments, decided row by row in the oracle's order): same answers as the oracle."""
    w, b = G.wb(name)
    m = eng.Model(w, b); net = O.Net(w, b)
    dom = G.domain(name)
    pa, ra = dom.columns.index("race"), dom.columns.index("age")
    grid = eng.Grid(dom.columns, dom.ranges, 6)
    first, n, S = len(grid) // 5, 14, 9
    boxes = grid.boxes(first, n); hb = boxes.cpu().numpy()
    samples = O.philox_samples(5, first, hb, S)
    want = O.partition_screen_ex(net, samples, hb, pa, ra, thr)
    got = m.partition_screen(boxes, pa, S=S, seed=5, first_index=first, ra_col=ra, ra_threshold=thr)
    for g_, w_, key in zip(got, want, ("ce_sample", "pa_pos", "pa_neg", "ra_dpos", "ra_dneg")):
        assert np.array_equal(g_.cpu().numpy(), w_), key
    got2 = m.partition_screen(boxes, pa, samples=samples, ra_col=ra, ra_threshold=thr)
    assert all(np.array_equal(a.cpu().numpy(), b_) for a, b_ in zip(got2, want))


@pytest.mark.parametrize("name", ["AC-1", "AC-2", "AC-10", "AC-4"])
def test_partition_screen_skip_mask_and_certified_boxes(eng, name):
    """skip: masked boxes report no counterexample, the others are untouched (strict and relaxed, every kernel family).
    certified(): a box the bounds kernel certifies one-signed -- widened on the relaxed column for a relaxed query --
    holds no counterexample in the unmasked screen either, so masking it changes nothing."""
    w, b = G.wb(name)
    m = eng.Model(w, b); net = O.Net(w, b)
    dom = G.domain(name)
    pa, ra, thr = dom.columns.index("race"), dom.columns.index("age"), 5
    grid = eng.Grid(dom.columns, dom.ranges, 6)
    n, S = (64, 40) if m.P > 4000 else (400, 60)
    # a window that straddles the first certified box of the grid's first 20 000, if there is one
    c0 = m.certified(grid.boxes(0, 20000)).cpu().numpy()
    first = max(0, int(np.flatnonzero(c0)[0]) - n // 2) if c0.any() else 2 * len(grid) // 3
    boxes = grid.boxes(first, n); hb = boxes.cpu().numpy()
    samples = O.philox_samples(3, first, hb, S)
    rng = np.random.default_rng(8)
    for relaxed in (False, True):
        kw = dict(ra_col=ra, ra_threshold=thr) if relaxed else {}
        want = O.partition_screen_ex(net, samples, hb, pa, ra if relaxed else -1, thr if relaxed else 0)
        nout = 5 if relaxed else 3
        mask = (rng.random(n) < 0.4).astype(np.uint8)
        verdict = torch.zeros((n,), dtype=torch.uint8, device="cuda")
        got = [t.cpu().numpy() for t in m.partition_screen(boxes, pa, S=S, seed=3, first_index=first, skip=mask, verdict=verdict, **kw)]
        none = (-1, -1, -1, 0, 0)
        for k in range(nout):
            assert np.array_equal(got[k], np.where(mask != 0, none[k], want[k])), (relaxed, k)
        assert np.array_equal((verdict.cpu().numpy() & eng.VERDICT_CE) != 0, (want[0] >= 0) & (mask == 0))
        cert = m.certified(boxes, **kw).cpu().numpy()
        assert not (want[0][cert != 0] >= 0).any()                  # a certified box has no counterexample
        got2 = [t.cpu().numpy() for t in m.partition_screen(boxes, pa, S=S, seed=3, first_index=first, skip=cert, **kw)]
        for k in range(nout):
            assert np.array_equal(got2[k], want[k]), (relaxed, k)
    if name in ("AC-1", "AC-2", "AC-10"):
        c = m.certified(boxes).cpu().numpy()                       # (these networks have certified boxes on this grid)
        assert c.any() and not c.all()
    with pytest.raises(Exception):
        m.partition_screen(boxes, pa, S=S, skip=np.zeros(n + 1, dtype=np.uint8))


@pytest.mark.parametrize("dims", [(13, 16, 8, 1), (13, 100, 100, 1), (13, 40, 24, 1)])
def test_relaxed_screen_on_networks_with_logits_at_the_rounding_level(eng, dims):
    """The interval pass that precedes a relaxed sweep must never set aside a sample whose sweep shows two classes, also
    where the logits are within float32 rounding of zero: random networks whose output bias cancels the typical logit
    (specialised kernel, tiled kernel, sample-per-lane kernel), against the oracle's full sweep."""
    rng = np.random.default_rng(sum(dims))
    dom = G.domain("AC-1")
    w = [(rng.standard_normal((dims[k], dims[k + 1])) / np.sqrt(dims[k])).astype(np.float32) for k in range(len(dims) - 1)]
    b = [(0.1 * rng.standard_normal(dims[k + 1])).astype(np.float32) for k in range(len(dims) - 1)]
    pa, ra, thr = dom.columns.index("race"), dom.columns.index("age"), 5
    grid = None
    for scale in (1.0, 1e-3):                       # second round: last layer scaled down, logits ~1e-3 of the first
        w[-1] = (w[-1] * np.float32(scale)).astype(np.float32)
        net0 = O.Net(w, b)
        m0 = eng.Model(w, b)
        grid = eng.Grid(dom.columns, dom.ranges, 6)
        first, n, S = len(grid) // 7, 96, 24
        boxes = grid.boxes(first, n); hb = boxes.cpu().numpy()
        samples = O.philox_samples(2, first, hb, S)
        y = O.forward_f32(net0, samples.reshape(-1, dims[0]).astype(np.float32))
        b[-1] = (b[-1] - np.float32(np.median(y))).astype(np.float32)      # half of the sampled logits on either side of 0
        net = O.Net(w, b); m = eng.Model(w, b)
        want = O.partition_screen_ex(net, samples, hb, pa, ra, thr)
        got = m.partition_screen(boxes, pa, S=S, seed=2, first_index=first, ra_col=ra, ra_threshold=thr)
        for g_, w_, key in zip(got, want, ("ce_sample", "pa_pos", "pa_neg", "ra_dpos", "ra_dneg")):
            assert np.array_equal(g_.cpu().numpy(), w_), (key, scale)
        assert (want[0] >= 0).any()
        del m0


@pytest.mark.parametrize("name", ["AC-1", "AC-5"])
def test_relaxed_screen_edge_cases(eng, name):
    """Threshold 0 (a relaxed query with nothing to relax: the strict pairs, relaxed output format), one sample per box,
    a sample count that is not a multiple of anything, a single box, and no box at all."""
    w, b = G.wb(name)
    m = eng.Model(w, b); net = O.Net(w, b)
    dom = G.domain(name)
    pa, ra = dom.columns.index("race"), dom.columns.index("age")
    grid = eng.Grid(dom.columns, dom.ranges, 6)
    first = len(grid) // 2
    for n, S, thr in ((40, 33, 0), (40, 1, 5), (1, 77, 3), (37, 5, 1)):
        boxes = grid.boxes(first, n); hb = boxes.cpu().numpy()
        samples = O.philox_samples(9, first, hb, S)
        want = O.partition_screen_ex(net, samples, hb, pa, ra, thr)
        got = m.partition_screen(boxes, pa, S=S, seed=9, first_index=first, ra_col=ra, ra_threshold=thr)
        for g_, w_, key in zip(got, want, ("ce_sample", "pa_pos", "pa_neg", "ra_dpos", "ra_dneg")):
            assert np.array_equal(g_.cpu().numpy(), w_), (key, n, S, thr)
        if thr == 0:
            strict = O.partition_screen(net, samples, hb, pa)
            assert np.array_equal(want[0], strict[0]) and not want[3].any() and not want[4].any()
    empty = m.partition_screen(grid.boxes(first, 0), pa, S=8, seed=1, ra_col=ra, ra_threshold=2)
    assert all(t.numel() == 0 for t in empty)


@pytest.mark.parametrize("name", ["AC-1", "AC-2", "AC-3"])
def test_reference_counterexample_csvs_on_the_device(eng, name):
    """The reference's shipped counterexample pairs (tests/golden/counterexamples.npz, see test_oracle_golden.py): device
    logits in float32 and float64 reproduce the printed sigmoid outputs and classes, and the strict partition screen finds
    a pair in the degenerate box of each first point (protected column over its range)."""
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "counterexamples.npz"))
    w, b = G.wb(name)
    m = eng.Model(w, b); dom = G.domain(name)
    X = z[f"{name}/x"]
    y32 = m.forward(X.astype(np.float32)).cpu().numpy().astype(np.float64)
    y64 = m.forward(X.astype(np.float64)).cpu().numpy()
    assert np.allclose(1.0 / (1.0 + np.exp(-y32)), z[f"{name}/output"], rtol=0, atol=5e-6)
    assert np.array_equal((y32 > 0).astype(np.uint8), z[f"{name}/prediction"])
    assert np.array_equal((y64 > 0).astype(np.uint8), z[f"{name}/prediction"])
    pa = dom.columns.index("sex")
    first = X[0::2]
    boxes = np.stack([first, first], axis=2).astype(np.int32)
    boxes[:, pa, 0], boxes[:, pa, 1] = dom.ranges["sex"]
    ce, vp, vn = [t.cpu().numpy() for t in m.partition_screen(boxes, pa, S=4, seed=1)]
    assert np.all(ce == 0) and np.all(vp != vn)                     # every sample of such a box is the point itself


def test_partition_screen_rejects_what_it_cannot_encode(eng):
    """A protected range wider than the key's fields is skipped and flagged, not silently mis-encoded; argument checks."""
    w, b = G.wb("GC-1")
    m = eng.Model(w, b); dom = G.domain("GC-1")
    col = dom.columns.index("credit_amount")                                   # range 0 .. 20000
    box = np.array([[dom.ranges[c][0], dom.ranges[c][1]] for c in dom.columns], dtype=np.int32)[None].repeat(3, 0)
    box[1, col] = [0, 70000]                                                   # 70 001 values: beyond 65 536
    box[2, col] = [-2**31, 2**31 - 1]                                          # the whole int32 range
    verdict = torch.zeros((3,), dtype=torch.uint8, device="cuda")
    ce = m.partition_screen(box, col, S=8, seed=1, verdict=verdict)[0].cpu().numpy()
    v = verdict.cpu().numpy()
    assert (v[0] & 8) == 0 and (v[1] & 8) and (v[2] & 8) and ce[1] == -1 and ce[2] == -1
    v2 = torch.zeros((3,), dtype=torch.uint8, device="cuda")
    m.partition_screen(box, col, S=8, seed=1, verdict=v2, ra_col=0, ra_threshold=2)     # relaxed: 256 values at most
    assert (v2.cpu().numpy() & 8).all()
    with pytest.raises(Exception):
        m.partition_screen(box, col, S=8, ra_col=col, ra_threshold=2)          # relaxed column == protected column
    with pytest.raises(Exception):
        m.partition_screen(box, col, S=8, ra_col=1, ra_threshold=500)


def test_hybrid_predict_matches_reference_functions(eng):
    """find_partition_result_for_point + hybrid_predict (src/AC/Verify-AC-experiment-new2.py:587-638) on the fixture the
    reference's own functions produced: which model answers each point, the debug counters, the class of every point."""
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "hybrid_case.npz"))
    dom = G.domain("AC-1")
    grid = eng.Grid(dom.columns, dom.ranges, int(z["T"]))
    orig, fair = eng.Model(*G.wb("AC-3")), eng.Model(*G.wb("AC-1"))
    logit, route, counters = eng.hybrid_predict(grid, orig, fair, z["X"], z["status"])
    logit = logit.cpu().numpy()
    assert np.array_equal(route.cpu().numpy(), z["route"])
    assert np.array_equal(counters.cpu().numpy(), z["counters"])
    assert np.array_equal(logit > 0, z["pred"] > 0.5)
    assert np.allclose(logit, z["logit"], rtol=2e-4, atol=2e-4)   # float32: numpy sums through BLAS, the oracle in index order
    lo = np.array([dom.ranges[c][0] for c in dom.columns]); hi = np.array([dom.ranges[c][1] for c in dom.columns])
    ko = np.array([dom.columns.index(k) for k in dom.ranges.keys()], dtype=np.int32)
    want, _ = O.hybrid_predict(O.Net(*G.wb("AC-3")), O.Net(*G.wb("AC-1")), lo, hi, int(z["T"]), ko, z["X"], z["status"])
    assert np.array_equal(logit, want)


def test_nearzero_diagnostic(eng):
    """fb_profile_nearzero equals the oracle on fixtures, and on the full src/AC grid (16 000 boxes x 1000 samples) no
    pre-activation is within 2^-40 of its terms: the activation counts do not depend on the summation order."""
    for name in ("AC-1", "GC-4", "BM-7"):
        m = eng.Model(*G.wb(name)); net = O.Net(*G.wb(name))
        cs = list(G.cases(name))[:2]
        for c in cs:
            got = m.nearzero(c["box"][None], samples=c["samples"][None])
            want = O.nearzero(net, c["samples"][None])
            assert got[0] == want[0] == 0 and got[1] == want[1], (name, got, want)
    m = eng.Model(*G.wb("AC-1")); dom = G.domain("AC-1")
    grid = eng.Grid(dom.columns, dom.ranges, 10)
    cnt, best = m.nearzero(grid.boxes(), S=1000, seed=2024)
    assert cnt == 0 and best > 2.0 ** -40, (cnt, best)


def test_prune_host_roundtrip_equals_device(eng):
    w, b = G.wb("AC-1")
    m = eng.Model(w, b)
    dom = G.domain("AC-1")
    grid = eng.Grid(dom.columns, dom.ranges, 10)
    boxes = grid.boxes(100, 5000)
    tol = _tol(eng, "AC-1")
    r = m.prune(boxes, S=64, seed=7, first_index=100, tol=tol)
    hb = boxes.cpu().pin_memory()
    out = m.prune_host(hb, S=64, seed=7, first_index=100, tol=tol)
    for k in ("cand", "b_dead", "dead", "sliver", "verdict", "counts"):
        assert np.array_equal(out[k].numpy(), r.numpy(k)), k


def test_prune_host_pipeline_slices_equal_device():
    """fb_prune_host cuts a batch into slices and overlaps upload / kernels / download on three streams; with the slice
    size forced down to 700 partitions (ragged last slice, injected samples too) every output must equal the
    device-resident run (subprocess: the slice size is read once per process)."""
    import subprocess
    import sys
    code = f"""
import numpy as np, sys, torch
sys.path.insert(0, {os.path.dirname(os.path.dirname(os.path.abspath(__file__)))!r})
from tests.common import Golden
from fairify_b200 import engine
from oracle import oracle as O
G = Golden()
for name in ("AC-1", "BM-4"):
    w, b = G.wb(name); m = engine.Model(w, b); dom = G.domain(name)
    grid = engine.Grid(dom.columns, dom.ranges, 10 if name == "AC-1" else dom.threshold)
    n = min(3005, len(grid)); boxes = grid.boxes(0, n)
    lo = np.array([dom.ranges[c][0] for c in dom.columns]); hi = np.array([dom.ranges[c][1] for c in dom.columns])
    tol = engine.sliver_tolerance(w, b, lo, hi)
    r = m.prune(boxes, S=48, seed=9, first_index=11, tol=tol, want=engine.ALL_OUTPUTS)
    out = m.prune_host(boxes.cpu().pin_memory(), S=48, seed=9, first_index=11, tol=tol, want=engine.ALL_OUTPUTS)
    for k in engine.ALL_OUTPUTS:
        assert np.array_equal(out[k].numpy(), r.numpy(k)), (name, k)
    samples = O.philox_samples(9, 11, boxes.cpu().numpy(), 48)
    out2 = m.prune_host(boxes.cpu(), samples=torch.as_tensor(samples), tol=tol)        # pageable host memory
    for k in ("dead", "verdict", "counts"):
        assert np.array_equal(out2[k].numpy(), r.numpy(k)), (name, k, "injected")
print("pipeline-ok")
"""
    env = dict(os.environ, FAIRIFY_B200_HOST_SLICE="700")
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=900)
    assert out.returncode == 0 and "pipeline-ok" in out.stdout, out.stdout + out.stderr


def test_empty_batch_and_bad_arguments(eng):
    w, b = G.wb("GC-3")
    m = eng.Model(w, b)
    r = m.prune(np.zeros((0, m.d, 2), dtype=np.int32), S=10)
    assert r.numpy("verdict").shape == (0,)
    with pytest.raises(Exception):
        m.prune(np.zeros((1, m.d, 2), dtype=np.int32), S=0)
    with pytest.raises(Exception):
        eng.Model([np.zeros((3, 2), np.float32)], [np.zeros(2, np.float32)])      # two logits
    with pytest.raises(Exception):
        eng.Model([np.full((3, 1), np.nan, np.float32)], [np.zeros(1, np.float32)])


def test_full_size_properties(eng):
    """At a size the oracle cannot finish quickly: size-independent properties of the 16 000-box src/AC grid,
    S = 1000.  (i) a bound-dead neuron is never seen active (soundness of the interval test),
    (ii) outward bounds enclose the nearest ones, (iii) certified partitions have no sampled counterexample,
    (iv) sharding invariance: any split of the index range gives the same bytes."""
    w, b = G.wb("AC-1")
    m = eng.Model(w, b)
    dom = G.domain("AC-1")
    grid = eng.Grid(dom.columns, dom.ranges, 10)
    boxes = grid.boxes()
    tol = _tol(eng, "AC-1")
    r = m.prune(boxes, S=1000, seed=2024, tol=tol, want=eng.ALL_OUTPUTS)
    counts = r.numpy("counts").view(np.uint16)
    b_dead, cand = r.bits("b_dead").astype(bool), r.bits("cand").astype(bool)
    assert np.all(counts[b_dead] == 0) and np.array_equal(cand, counts == 0)
    assert np.all(r.numpy("lb_rd") <= r.numpy("ws_lb")) and np.all(r.numpy("ub_ru") >= r.numpy("ws_ub"))
    assert np.all(counts[:, -1] == 1000)            # the logit is never exactly 0
    le0 = r.numpy("ws_ub") <= 0
    expect = cand & le0
    expect[:, -1] = cand[:, -1]
    for l in range(m.L):                            # keep-one fix, utils/prune.py:689-691
        sl = slice(int(m.boff[l]), int(m.boff[l + 1]))
        alldead = expect[:, sl].all(axis=1)
        expect[alldead, int(m.boff[l])] = False
    assert np.array_equal(b_dead, expect)
    pa_col = dom.columns.index("sex")
    verdict = r.t["verdict"].clone()
    ce = m.partition_screen(boxes, pa_col, S=1000, seed=2024, verdict=verdict)[0].cpu().numpy()
    v = verdict.cpu().numpy()
    assert not np.any((v & 1).astype(bool) & (ce >= 0))
    # sharding invariance (what the multi-GPU path relies on)
    cut = 7001
    r1 = m.prune(boxes[:cut], S=1000, seed=2024, first_index=0, tol=tol)
    r2 = m.prune(boxes[cut:], S=1000, seed=2024, first_index=cut, tol=tol)
    for k in ("cand", "b_dead", "verdict", "counts"):
        assert np.array_equal(np.concatenate([r1.numpy(k), r2.numpy(k)]), r.numpy(k)), k


def _two_rank_worker(rank, world, port, ret):
    import torch.distributed as dist
    from fairify_b200 import engine
    from fairify_b200.dist import prune_grid_sharded
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.cuda.set_device(rank % torch.cuda.device_count())
        w, b = G.wb("GC-1")
        m = engine.Model(w, b)
        dom = G.domain("GC-1")
        grid = engine.Grid(dom.columns, dom.ranges, 10)            # stress/GC grid, 18 009 boxes (odd: uneven shards)

        class CpuGather:                                           # gloo gathers host tensors
            pass
        import fairify_b200.dist as D
        first, count = D.shard_range(len(grid), rank, world)
        pa = dom.columns.index(dom.pa[0])
        bx = grid.boxes(first, count)
        r = m.prune(bx, S=64, seed=5, first_index=first, want=("record",))
        m.partition_screen(bx, pa, S=64, seed=5, first_index=first, record=r.t["record"])
        full = D.allgather_records(r.t["record"].cpu(), len(grid))
        v, d, c, ce = D.unpack_records(full, m.NW)
        one = m.prune(grid.boxes(), S=64, seed=5, first_index=0, want=("cand", "dead", "verdict"))
        one_ce = m.partition_screen(grid.boxes(), pa, S=64, seed=5, first_index=0, verdict=one.t["verdict"])[0]
        ok = (np.array_equal(v.numpy(), one.numpy("verdict")) and np.array_equal(d.numpy(), one.numpy("dead"))
              and np.array_equal(c.numpy(), one.numpy("cand")) and np.array_equal(ce.numpy(), one_ce.cpu().numpy()))
        ret[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


def test_two_rank_sharded_run_equals_single_rank(eng):
    """two processes, one contiguous shard each (both on the visible GPU(s)), records gathered with gloo:
    byte-identical to the single-process run of the whole grid."""
    import socket
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
    mgr = mp.Manager(); ret = mgr.dict()
    mp.spawn(_two_rank_worker, args=(2, port, ret), nprocs=2, join=True)
    assert ret.get(0) and ret.get(1), dict(ret)


@pytest.mark.parametrize("name", MODELS)
def test_heuristic_prune_batched(eng, name):
    """fb_heuristic on all golden cases of a model at once, thresholds 5/20/50/100 (utils/prune.py:862-939)."""
    w, b = G.wb(name)
    m = eng.Model(w, b)
    cs = list(G.cases(name))
    lb = np.stack([c["ws_lb"] for c in cs]); ub = np.stack([c["ws_ub"] for c in cs])
    cand = eng.pack_bits(np.stack([c["cand"] for c in cs])); bc = eng.pack_bits(np.stack([c["b_cand"] for c in cs]))
    bd = eng.pack_bits(np.stack([c["b_dead"] for c in cs]))
    for pct in (5, 20, 50, 100):
        new, merged = m.heuristic(lb, ub, cand, bc, bd, pct)
        new = eng.unpack_bits(new.cpu().numpy().view(np.uint32), m.N); merged = eng.unpack_bits(merged.cpu().numpy().view(np.uint32), m.N)
        for k, c in enumerate(cs):
            assert np.array_equal(new[k], c[f"h{pct}_new"]), f"{c['tag']} pct={pct}"
            assert np.array_equal(merged[k], c[f"h{pct}_merged"]), f"{c['tag']} pct={pct}"


@pytest.mark.parametrize("tag", ["AC_T30", "AC_T12"])
def test_grid_locate_and_density(eng, tag):
    """fb_grid_locate == oracle == the reference's p_list_density; the drop-in p_list_density goes through it."""
    import pandas as pd
    from fairify_b200.domains import DOMAINS
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "density.npz"))
    D = DOMAINS["AC"]; T = int(z[f"{tag}/T"]); X = z[f"{tag}/X"]
    grid = eng.Grid(D.columns, D.ranges, T)
    idx, dens = grid.locate(X.astype(np.float64), want_density=True)
    lo = [D.ranges[c][0] for c in D.columns]; hi = [D.ranges[c][1] for c in D.columns]
    ko = [D.columns.index(k) for k in D.ranges.keys()]
    assert np.array_equal(idx.cpu().numpy(), O.grid_locate(lo, hi, T, ko, X))
    assert np.array_equal(dens.cpu().numpy() / X.shape[0], z[f"{tag}/density"])
    from utils.input_partition import p_list_density, partition, partitioned_ranges
    rd = D.copy_ranges()
    p_list = partitioned_ranges(rd.keys(), D.pa, partition(rd, T), rd)
    df = pd.DataFrame(np.column_stack([X, np.zeros(len(X), dtype=np.int64)]), columns=D.columns + [D.label])
    assert p_list_density(D.copy_ranges(), p_list, df) == list(z[f"{tag}/density"])


def test_result_persistence_roundtrip(eng, tmp_path):
    from fairify_b200 import persist
    w, b = G.wb("GC-4")
    m = eng.Model(w, b)
    dom = G.domain("GC-4")
    grid = eng.Grid(dom.columns, dom.ranges, 10)
    first, n = 1000, 2000
    r = m.prune(grid.boxes(first, n), S=64, seed=9, first_index=first)
    path = str(tmp_path / "gc4.npz")
    persist.save(path, m, grid, 64, 9, first, r)
    back = persist.load(path, m, grid, 64, 9, first, n)
    for k in ("cand", "b_dead", "dead", "verdict", "counts"):
        assert np.array_equal(back[k], r.numpy(k))
    with pytest.raises(ValueError):
        persist.load(path, m, grid, 64, 10, first, n)                    # other seed
    with pytest.raises(ValueError):
        persist.load(path, eng.Model(*G.wb("GC-3")), grid, 64, 9, first, n)   # other model
