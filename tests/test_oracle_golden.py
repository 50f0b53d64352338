"""CPU: the oracle (oracle/fairify_oracle.c) against the golden vectors the UNMODIFIED reference produced
(tests/golden/make_golden.py).  This is what pins the oracle; the GPU tests then compare against it."""
import hashlib
import os

import numpy as np
import pytest

from oracle import oracle as O
from tests.common import Golden

G = Golden()
MODELS = G.models


@pytest.mark.parametrize("tag", ["AC_T10", "GC_T100", "BM_T100", "GC_T10", "CP_T5", "AC_T6_age30_35",
                                 "GC_T10_purpose7", "BM_T10_small"])
def test_grid_matches_reference(tag):
    """partition + partitioned_ranges (utils/input_partition.py:17-76): count, every box (sha256), sampled boxes."""
    g = G.grids
    lo, hi, T, ko = g[f"{tag}/lo"], g[f"{tag}/hi"], int(g[f"{tag}/T"]), g[f"{tag}/key_order"]
    n, _ = O.grid_count(lo, hi, T)
    assert n == int(g[f"{tag}/n"])
    boxes = O.grid_boxes(lo, hi, T, ko, 0, n)
    assert hashlib.sha256(boxes.tobytes()).digest() == g[f"{tag}/sha256"].tobytes()
    for k, idx in enumerate(g[f"{tag}/idx"]):
        assert np.array_equal(O.grid_boxes(lo, hi, T, ko, int(idx), 1)[0], g[f"{tag}/boxes"][k])


@pytest.mark.parametrize("name", MODELS)
def test_bounds_bit_exact(name):
    """neuron_bounds (utils/prune.py:105-164): all four bound arrays bit-for-bit."""
    net = O.Net(*G.wb(name))
    for c in G.cases(name):
        ws_lb, ws_ub, pl_lb, pl_ub = O.ibp(net, c["box"][None])
        for got, key in ((ws_lb, "ws_lb"), (ws_ub, "ws_ub"), (pl_lb, "pl_lb"), (pl_ub, "pl_ub")):
            assert np.array_equal(got[0], c[key]), f"{c['tag']} {key}"


@pytest.mark.parametrize("name", MODELS)
def test_directed_bounds_enclose_reference(name):
    """outward-rounded enclosure contains the reference bounds and is within 1e-9 relative of them."""
    net = O.Net(*G.wb(name))
    for c in G.cases(name):
        lb_rd, ub_ru = O.ibp_directed(net, c["box"][None])
        assert np.all(lb_rd[0] <= c["ws_lb"]) and np.all(ub_ru[0] >= c["ws_ub"]), c["tag"]
        scale = np.maximum(1.0, np.maximum(np.abs(c["ws_lb"]), np.abs(c["ws_ub"])))
        assert np.all((c["ws_lb"] - lb_rd[0]) <= 1e-9 * scale) and np.all((ub_ru[0] - c["ws_ub"]) <= 1e-9 * scale)


@pytest.mark.parametrize("name", MODELS)
def test_profile_and_masks(name):
    """candidate_dead_nodes (:168-192), dead_node_from_bound (:226-251) + keep-one fix, compression_ratio."""
    net = O.Net(*G.wb(name))
    dom = G.domain(name)
    lo = np.array([dom.ranges[c][0] for c in dom.columns]); hi = np.array([dom.ranges[c][1] for c in dom.columns])
    tol = O.sliver_tolerance(net, lo, hi)
    for c in G.cases(name):
        counts = O.profile(net, c["samples"][None])
        S = int(c["S"])
        assert np.array_equal(counts[0] == 0, c["cand"].astype(bool)), c["tag"]
        assert np.array_equal(counts[0] / S, c["pos_prob"]), c["tag"]
        _, ws_ub, pl_lb, pl_ub = O.ibp(net, c["box"][None])
        m = O.masks(net, counts, ws_ub, pl_lb, pl_ub, tol)
        assert np.array_equal(m["cand"][0], c["cand"])
        assert np.array_equal(m["b_dead"][0], c["b_dead"]), c["tag"]
        assert np.array_equal(m["b_cand"][0], c["b_cand"]), c["tag"]
        assert O.compression_ratio(m["b_dead"][0]) == float(c["b_compression"])
        assert not m["sliver"].any()
        if "dead" in c:       # full sound_prune* run including the Z3 singular_verification
            assert np.array_equal(m["s_dead"][0], c["s_dead"]), c["tag"]
            assert np.array_equal(m["dead"][0], c["dead"]), c["tag"]
            assert np.array_equal(m["s_cand"][0], c["s_cand"]), c["tag"]
            assert O.compression_ratio(m["dead"][0]) == float(c["st_compression"])
            assert O.compression_ratio(m["s_dead"][0]) == float(c["s_compression"])


@pytest.mark.parametrize("name", G.singular_models)
def test_singular_verification_closed_form(name):
    """sound_prune* INCLUDING the Z3 singular_verification (utils/prune.py:276-644) on every model with a hidden layer
    of equal input and output width (plus three funnels): the reference bounds a layer's inputs with that layer's own
    post-activation bounds (:266-273 called with layer_index, :302-345), so it finds neurons dead whose interval upper
    bound is far above zero.  The closed form must give the reference's s_dead / s_candidates / dead_nodes exactly."""
    net = O.Net(*G.wb(name))
    dom = G.domain(name)
    lo = np.array([dom.ranges[c][0] for c in dom.columns]); hi = np.array([dom.ranges[c][1] for c in dom.columns])
    tol = O.sliver_tolerance(net, lo, hi)
    hits = 0
    for c in G.singular_cases(name):
        counts = O.profile(net, c["samples"][None])
        _, ws_ub, pl_lb, pl_ub = O.ibp(net, c["box"][None])
        assert np.array_equal(ws_ub[0], c["ws_ub"]) and np.array_equal(pl_lb[0], c["pl_lb"]) and np.array_equal(pl_ub[0], c["pl_ub"])
        m = O.masks(net, counts, ws_ub, pl_lb, pl_ub, tol)
        for k in ("cand", "b_dead", "s_dead", "s_cand", "dead"):
            assert np.array_equal(m[k][0], c[k]), f"{c['tag']} {k}"
        assert not m["sliver"].any(), c["tag"]
        assert O.compression_ratio(m["dead"][0]) == float(c["st_compression"])
        assert O.compression_ratio(m["s_dead"][0]) == float(c["s_compression"])
        hits += int(c["s_dead"].sum())
        if hits:          # a neuron Z3 found dead although its interval bound is positive
            sd = c["s_dead"].astype(bool)
            assert np.all(c["ws_ub"][sd] > 0)
    if name in ("AC-4", "AC-8", "AC-10", "AC-12", "BM-6", "BM-7"):
        assert hits > 0, "the fixtures of this model hold partitions where the reference's Z3 pass finds a neuron"


def test_singular_verification_ac10_partition_13529():
    """The partition the round-1 review named: src/AC grid index 13529, layer 1 neuron 1, ws_ub = 14.47, s_dead = 1."""
    net = O.Net(*G.wb("AC-10"))
    c = next(c for c in G.singular_cases("AC-10") if int(c["index"]) == 13529)
    j = int(net.boff[1]) + 1
    assert c["s_dead"][j] == 1 and abs(c["ws_ub"][j] - 14.466598695433476) < 1e-9
    _, ws_ub, pl_lb, pl_ub = O.ibp(net, c["box"][None])
    m = O.masks(net, O.profile(net, c["samples"][None]), ws_ub, pl_lb, pl_ub, None)
    assert m["s_dead"][0][j] == 1 and m["q"][0][j] < -10.0 and m["dead"][0][j] == 1


def test_known_answers_from_the_reference_notebook():
    """reproduce.ipynb: B_compression plateaus at 0.7 (GC-3) and 0.6364 (GC-4) on the src/GC grid."""
    vals = {}
    for name in ("GC-3", "GC-4"):
        vals[name] = [round(float(c["b_compression"]), 4) for c in G.cases(name)]
    assert 0.7 in vals["GC-3"] and 0.6364 in vals["GC-4"]


@pytest.mark.parametrize("name", MODELS)
def test_heuristic_prune(name):
    """heuristic_prune (:862-939) for the thresholds the drivers use (5, 20, 50, 100)."""
    net = O.Net(*G.wb(name))
    for c in G.cases(name):
        for pct in (5, 20, 50, 100):
            new, merged = O.heuristic_prune(net, c["ws_lb"], c["ws_ub"], c["cand"], c["b_cand"], c["b_dead"], pct)
            assert np.array_equal(new, c[f"h{pct}_new"]), f"{c['tag']} pct={pct}"
            assert np.array_equal(merged, c[f"h{pct}_merged"]), f"{c['tag']} pct={pct}"


def test_numpy_statistics_restatement():
    """np.mean / np.median / np.percentile as restated in C, bit-exact on random vectors of many lengths."""
    import ctypes as C
    s = G.stats
    lib = O.lib()
    off = 0
    for n, want in zip(s["lens"], s["res"]):
        v = np.ascontiguousarray(s["data"][off:off + n]); off += n
        srt = np.sort(v)
        p = v.ctypes.data_as(C.c_void_p); ps = srt.ctypes.data_as(C.c_void_p)
        got = [lib.fo_np_mean(p, int(n)), lib.fo_np_median_sorted(ps, int(n))]
        for pct in (0, 5, 20, 50, 80, 95, 100):
            got.append(lib.fo_np_percentile_sorted(ps, int(n), C.c_double(pct)))
        assert np.array_equal(np.array(got), want), f"n={n}"


@pytest.mark.parametrize("name", MODELS)
def test_logits(name):
    """net (utils/<M>-Model-Functions.py): sign always, value to 1e-12 (f64) / 1e-5 (f32) relative --
    the reference sums through BLAS in an unspecified order, the oracle in index order."""
    net = O.Net(*G.wb(name))
    for c in G.cases(name):
        rows = c["samples"][:64]
        y64 = O.forward_f64(net, rows.astype(np.float64))
        y32 = O.forward_f32(net, rows.astype(np.float32))
        assert np.array_equal(y64 > 0, c["logit64"] > 0) and np.array_equal(y32 > 0, c["logit32"] > 0)
        assert np.allclose(y64, c["logit64"], rtol=1e-12, atol=1e-12)
        assert np.allclose(y32, c["logit32"], rtol=1e-4, atol=1e-4)


def test_philox_reference_vector():
    """Philox4x32-10 known-answer test (Random123 kat_vectors: counter = key = 0 and all-ones)."""
    import ctypes as C
    # exercised through fo_philox_samples: range 2^31 on feature 0..3 exposes the raw words' top bits
    boxes = np.array([[[0, 2 ** 31 - 1]] * 4], dtype=np.int32)
    s = O.philox_samples(0, 0, boxes, 1)[0, 0]
    want = np.array([0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8], dtype=np.uint64) >> np.uint64(1)
    assert np.array_equal(s.astype(np.uint64), want)


def test_philox_samples_in_range_and_uniform():
    boxes = np.array([[[10, 19], [0, 0], [-3, 1], [0, 20000], [5, 6]]], dtype=np.int32)
    s = O.philox_samples(1234, 7, boxes, 4000)[0]
    assert np.all(s >= boxes[0, :, 0]) and np.all(s <= boxes[0, :, 1])
    for i in (0, 2, 4):
        vals, cnt = np.unique(s[:, i], return_counts=True)
        assert len(vals) == boxes[0, i, 1] - boxes[0, i, 0] + 1
        assert cnt.min() > 0.7 * 4000 / len(vals)


def _philox_py(c, k):
    c = list(c); k = list(k); M = 0xFFFFFFFF
    for _ in range(10):
        p0 = 0xD2511F53 * c[0]; p1 = 0xCD9E8D57 * c[2]
        c = [(p1 >> 32) ^ c[1] ^ k[0], p1 & M, (p0 >> 32) ^ c[3] ^ k[1], p0 & M]
        k = [(k[0] + 0x9E3779B9) & M, (k[1] + 0xBB67AE85) & M]
    return c


def _draw_py(seed, pidx, s, box):
    """Pure-Python restatement of the batched bounded-integer draw (fb_common.cuh / fo_sample_row)."""
    widest = max(int(hi) - int(lo) + 1 for lo, hi in box)
    G = 4 if widest <= 16 else (2 if widest <= 256 else 1)
    out = []; word = 0; r = None
    for i, (lo, hi) in enumerate(box):
        if i % G == 0:
            q = i // G
            if q % 4 == 0:
                r = _philox_py([s, pidx & 0xFFFFFFFF, pidx >> 32, q // 4], [seed & 0xFFFFFFFF, seed >> 32])
            word = r[q % 4]
        prod = word * (int(hi) - int(lo) + 1)
        out.append(int(lo) + (prod >> 32)); word = prod & 0xFFFFFFFF
    return out


@pytest.mark.parametrize("widths", [[4, 3, 1, 2, 4, 4, 2, 16, 5, 1, 3, 2, 4], [100, 7, 2, 256, 30, 1, 9, 64, 3],
                                    [100, 20001, 5, 2, 80, 3], [10] * 20])
def test_batched_draw_matches_python_restatement(widths):
    """All three draw classes (4, 2, 1 features per Philox word) against an independent pure-Python restatement."""
    rng = np.random.RandomState(5)
    lo = rng.randint(-50, 50, size=len(widths))
    boxes = np.stack([lo, lo + np.array(widths) - 1], axis=1).astype(np.int32)[None]
    got = O.philox_samples(0x1234567890ABCDEF, (1 << 33) + 11, boxes, 40)[0]
    for s in range(40):
        assert got[s].tolist() == _draw_py(0x1234567890ABCDEF, (1 << 33) + 11, s, boxes[0])


@pytest.mark.parametrize("widths", [[4, 3, 16, 2, 5, 4, 4, 4, 7], [100, 7, 256, 2, 30, 30]])
def test_batched_draw_is_uniform_and_independent(widths):
    """Values drawn from one shared Philox word are uniform and pairwise independent (chi-square, 6 sigma)."""
    S = 200000
    boxes = np.stack([np.zeros(len(widths)), np.array(widths) - 1], axis=1).astype(np.int32)[None]
    s = O.philox_samples(99, 3, boxes, S)[0]
    for i, w in enumerate(widths):
        cnt = np.bincount(s[:, i], minlength=w)
        chi2 = ((cnt - S / w) ** 2 / (S / w)).sum()
        assert chi2 < (w - 1) + 6 * np.sqrt(2 * max(w - 1, 1)), (i, chi2)
    for i in range(len(widths) - 1):                    # neighbours (share a word whenever they are in one group)
        wa, wb = widths[i], widths[i + 1]
        joint = np.bincount(s[:, i] * wb + s[:, i + 1], minlength=wa * wb)
        dof = wa * wb - 1
        chi2 = ((joint - S / (wa * wb)) ** 2 / (S / (wa * wb))).sum()
        assert chi2 < dof + 6 * np.sqrt(2 * dof), (i, chi2)


@pytest.mark.parametrize("tag", ["AC_T30", "AC_T12"])
def test_point_to_partition_density(tag):
    """p_list_density (utils/input_partition.py:184-218): rows per partition via index arithmetic == the reference's scan."""
    from fairify_b200.domains import DOMAINS
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "density.npz"))
    D = DOMAINS["AC"]
    lo = [D.ranges[c][0] for c in D.columns]; hi = [D.ranges[c][1] for c in D.columns]
    ko = [D.columns.index(k) for k in D.ranges.keys()]
    X = z[f"{tag}/X"]; n = int(z[f"{tag}/n"])
    idx = O.grid_locate(lo, hi, int(z[f"{tag}/T"]), ko, X)
    dens = np.bincount(idx[idx >= 0], minlength=n) / X.shape[0]
    assert np.array_equal(dens, z[f"{tag}/density"])
    # located boxes really contain their rows
    boxes = O.grid_boxes(lo, hi, int(z[f"{tag}/T"]), ko, 0, n)
    inside = idx >= 0
    b = boxes[idx[inside]]
    assert np.all((X[inside] >= b[:, :, 0]) & (X[inside] <= b[:, :, 1]))
    assert not np.any(np.all((X[~inside][:, None, :] >= boxes[None, :, :, 0]) & (X[~inside][:, None, :] <= boxes[None, :, :, 1]), axis=2))
    # non-integer points between two chunks belong to no partition
    Xf = X[inside][:5].astype(np.float64); Xf[:, 0] = 39.5 if int(z[f"{tag}/T"]) == 30 else 21.5
    assert np.all(O.grid_locate(lo, hi, int(z[f"{tag}/T"]), ko, Xf) == -1)


def test_constructed_sliver_case_is_flagged():
    """tests/golden/sliver_case.npz: the reference (with Z3) on a network built so that ws_ub of layer-0 neuron 0 is
    2^-50 -- inside the rounding tolerance.  The oracle must flag exactly that neuron and otherwise agree."""
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "sliver_case.npz"))
    w = [z[f"w{l}"] for l in range(3)]; b = [z[f"b{l}"] for l in range(3)]
    net = O.Net(w, b)
    from tests.common import reference_samples
    box = z["box"]
    samples = reference_samples(z["seed"], box, z["S"])
    assert hashlib.sha256(samples.tobytes()).digest() == z["sim_sha256"].tobytes()
    tol = O.sliver_tolerance(net, box[:, 0], box[:, 1])
    _, ws_ub, pl_lb, pl_ub = O.ibp(net, box[None])
    assert np.array_equal(ws_ub[0], z["ws_ub"]) and ws_ub[0][0] == 2.0 ** -50
    m = O.masks(net, O.profile(net, samples[None]), ws_ub, pl_lb, pl_ub, tol)
    assert m["sliver"][0][0] == 1 and m["sliver"][0].sum() == 1
    for k in ("cand", "b_dead", "s_dead", "s_cand", "dead"):
        assert np.array_equal(m[k][0], z[k]), k


def test_hybrid_predict_matches_reference_functions():
    """tests/golden/hybrid_case.npz: find_partition_result_for_point + hybrid_predict of
    src/AC/Verify-AC-experiment-new2.py:587-638 on 300 points and a random verdict table (original AC-3, fairer AC-1)."""
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "hybrid_case.npz"))
    dom = G.domain("AC-1")
    lo = np.array([dom.ranges[c][0] for c in dom.columns]); hi = np.array([dom.ranges[c][1] for c in dom.columns])
    ko = np.array([dom.columns.index(k) for k in dom.ranges.keys()], dtype=np.int32)
    orig, fair = O.Net(*G.wb("AC-3")), O.Net(*G.wb("AC-1"))
    logit, route = O.hybrid_predict(orig, fair, lo, hi, int(z["T"]), ko, z["X"], z["status"])
    assert np.array_equal(route, z["route"])
    assert np.array_equal(np.bincount(route, minlength=4), z["counters"])
    assert np.array_equal(logit > 0, z["logit"] > 0) and np.allclose(logit, z["logit"], rtol=2e-4, atol=2e-4)   # float32: numpy sums through BLAS, the oracle in index order
    assert np.array_equal(logit > 0, z["pred"] > 0.5)


@pytest.mark.parametrize("name", ["AC-1", "AC-2", "AC-3"])
def test_reference_counterexample_csvs_are_reproduced(name):
    """Known answers shipped with the reference (experimentData/task5/AC/counterexamples-fairify-AC-*.csv, re-encoded by
    make_golden.py counterexamples): `output` = sigmoid(net(x)) of the float32-cast point as the reference's driver
    printed it (float32 arithmetic, pruned network), `prediction` = its class; consecutive rows are a counterexample pair."""
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "counterexamples.npz"))
    net = O.Net(*G.wb(name))
    X = z[f"{name}/x"]
    y = O.forward_f32(net, X.astype(np.float32)).astype(np.float64)
    assert np.allclose(1.0 / (1.0 + np.exp(-y)), z[f"{name}/output"], rtol=0, atol=5e-6)
    assert np.array_equal((y > 0).astype(np.uint8), z[f"{name}/prediction"])
    y64 = O.forward_f64(net, X.astype(np.float64))
    assert np.array_equal((y64 > 0).astype(np.uint8), z[f"{name}/prediction"])
    pred = z[f"{name}/prediction"]
    assert np.all(pred[0::2] != pred[1::2])                               # every pair flips the class
    diff = X[0::2] != X[1::2]
    assert np.all(diff.sum(axis=1) == 1) and np.all(diff[:, G.domain(name).columns.index("sex")])   # in the protected column only


def test_relaxed_pairs_are_the_ones_the_reference_property_admits():
    """tests/golden/relaxed_case.npz: 720 concrete pairs judged by the property relaxed/AC/Verify-AC.py:122-132 builds
    with the reference's own in_const_adult / in_const_diff_adult / in_const_domain_adult (Z3, make_golden.py relaxed).
    The set the screen enumerates -- and test_relaxed_screen_restatement brute-forces -- must be that set: first point
    in the box, second point equal to it except for another protected value inside the box's protected range and a
    relaxed attribute at most `thr` away (not confined to the box)."""
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "relaxed_case.npz"))
    box, x, x2 = z["box"].astype(np.int64), z["x"], z["x2"]
    pa, ra, thr = int(z["pa"]), int(z["ra"]), int(z["thr"])
    other = np.ones(x.shape[1], dtype=bool); other[[pa, ra]] = False
    rule = (np.all((x >= box[:, :, 0]) & (x <= box[:, :, 1]), axis=1)
            & (x2[:, pa] >= box[:, pa, 0]) & (x2[:, pa] <= box[:, pa, 1]) & (x2[:, pa] != x[:, pa])
            & (np.abs(x2[:, ra] - x[:, ra]) <= thr)
            & np.all(x2[:, other] == x[:, other], axis=1))
    assert np.array_equal(rule, z["admissible"] != 0)
    assert 100 < rule.sum() < len(rule) - 100                       # both answers well represented
    beyond = rule & ((x2[:, ra] < box[:, ra, 0]) | (x2[:, ra] > box[:, ra, 1]))
    assert beyond.any()                                             # second points outside the box on the relaxed attribute


def test_relaxed_screen_restatement():
    """The relaxed screen against a brute-force enumeration of the property the relaxed drivers give Z3
    (relaxed/AC/Verify-AC.py:128-132): x in the box, x' equal to x except on the protected column (different value, inside
    the box's range) and on the relaxed column (|difference| <= threshold), classes strictly different."""
    name = "AC-3"
    net = O.Net(*G.wb(name)); dom = G.domain(name)
    pa, ra, thr = dom.columns.index("race"), dom.columns.index("age"), 5
    c = next(iter(G.cases(name)))
    rng = np.random.RandomState(3)
    boxes = np.repeat(c["box"][None], 6, axis=0).copy()
    boxes[:, 0, 0] = [10, 30, 50, 70, 90, 95]; boxes[:, 0, 1] = boxes[:, 0, 0] + 5
    samples = rng.randint(boxes[:, None, :, 0], boxes[:, None, :, 1] + 1, size=(6, 40, dom.d)).astype(np.int32)
    ce, vp, vn, dp, dn = O.partition_screen_ex(net, samples, boxes, pa, ra, thr)
    strict = O.partition_screen(net, samples, boxes, pa)
    found_relaxed_only = 0
    for p in range(6):
        lo, hi = boxes[p, pa]
        first = -1
        for s in range(40):
            x = samples[p, s].astype(np.float32)
            base = []
            for v in range(lo, hi + 1):
                xv = x.copy(); xv[pa] = v
                base.append(O.forward_f32(net, xv[None])[0])
            exists = False
            for i, v in enumerate(range(lo, hi + 1)):
                for v2 in range(lo, hi + 1):
                    if v2 == v:
                        continue
                    for d in range(-thr, thr + 1):
                        x2 = x.copy(); x2[pa] = v2; x2[ra] = x[ra] + d
                        y2 = O.forward_f32(net, x2[None])[0]
                        if (base[i] > 0 and y2 < 0) or (base[i] < 0 and y2 > 0):
                            exists = True
            if exists:
                first = s
                break
        assert ce[p] == first, (p, ce[p], first)
        if strict[0][p] >= 0 and strict[0][p] == ce[p]:
            assert (vp[p], vn[p], dp[p], dn[p]) == (strict[1][p], strict[2][p], 0, 0)
        if ce[p] >= 0:
            x = samples[p, ce[p]].astype(np.float32)
            xp = x.copy(); xp[pa] = vp[p]; xp[ra] += dp[p]
            xn = x.copy(); xn[pa] = vn[p]; xn[ra] += dn[p]
            assert O.forward_f32(net, xp[None])[0] > 0 > O.forward_f32(net, xn[None])[0]
            assert vp[p] != vn[p] and (dp[p] == 0 or dn[p] == 0) and max(abs(dp[p]), abs(dn[p])) <= thr
            found_relaxed_only += int(strict[0][p] != ce[p])
    assert np.array_equal(O.partition_screen_ex(net, samples, boxes, pa, -1, 0)[0], strict[0])


def test_no_preactivation_sign_depends_on_the_summation_order():
    """On every fixture: the smallest |pre-activation| relative to the size of its terms is far above the rounding error a
    different summation order (numpy's BLAS, the reference) could introduce, so the activation counts are order-proof."""
    worst = 1.0
    for name in MODELS:
        net = O.Net(*G.wb(name))
        for c in G.cases(name):
            cnt, best = O.nearzero(net, c["samples"][None])
            assert cnt == 0, (c["tag"], cnt, best)
            worst = min(worst, best)
    assert worst > 2.0 ** -40
