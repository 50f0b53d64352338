"""The drop-in `utils` package (same names / signatures / return structures as the reference's utils/*):
host-only pieces on CPU, the GPU-backed entry points under the gpu marker."""
import hashlib
import importlib
import os

import numpy as np
import pandas as pd
import pytest

from tests.common import Golden

G = Golden()


def _df(dom):
    return pd.DataFrame(np.zeros((2, dom.d + 1)), columns=dom.columns + [dom.label])


def _box_dict(dom, box):
    return {c: [int(box[i, 0]), int(box[i, 1])] for i, c in enumerate(dom.columns)}


def _lists(model_dims, flat, cast=float):
    off = np.concatenate([[0], np.cumsum(model_dims[1:])])
    return [[cast(v) for v in flat[off[l]:off[l + 1]]] for l in range(len(model_dims) - 1)]


# ---------------------------------------------------------------------------------------------------
# CPU
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("tag", ["AC_T10", "GC_T100", "BM_T100", "GC_T10", "CP_T5", "BM_T10_small"])
def test_partition_grid_matches_reference(tag):
    from utils.input_partition import partition, partitioned_ranges
    from fairify_b200.domains import DOMAINS
    g = G.grids
    dom = DOMAINS[tag.split("_")[0]]
    rd = dom.copy_ranges()
    for c, a, b in zip(dom.columns, g[f"{tag}/lo"], g[f"{tag}/hi"]):
        rd[c] = [int(a), int(b)]
    p_dict = partition(rd, int(g[f"{tag}/T"]))
    p_list = partitioned_ranges(rd.keys(), dom.pa, p_dict, rd)
    assert len(p_list) == int(g[f"{tag}/n"])
    boxes = np.array([[p[c] for c in dom.columns] for p in p_list], dtype=np.int32)
    assert hashlib.sha256(boxes.tobytes()).digest() == g[f"{tag}/sha256"].tobytes()
    assert list(p_list[0].keys())[-len(p_dict):] == list(p_dict.keys())      # un-split attributes come first
    assert p_list[5 % len(p_list)].index == 5 % len(p_list)
    import random
    random.shuffle(p_list)                                                     # the drivers shuffle the list


def test_partition_list_position_survives_shuffle():
    """PartitionList.position (what lets sound_prune* batch the list's next entries) follows the list's current order."""
    import random
    from utils.input_partition import PartitionBox, partition, partitioned_ranges
    from fairify_b200.domains import DOMAINS
    dom = DOMAINS["GC"]
    rd = dom.copy_ranges()
    p_list = partitioned_ranges(rd.keys(), dom.pa, partition(rd, dom.threshold), rd)
    assert all(p_list.position(p) == i for i, p in enumerate(p_list))
    random.Random(1).shuffle(p_list)
    assert [p_list.position(p) for p in p_list[:50]] == list(range(50))
    assert sorted(p.index for p in p_list) == list(range(len(p_list)))
    assert p_list.position(PartitionBox(p_list[0])) is None                    # an equal copy is not a member
    p_list[3]["month"][0] = -1                                                 # every box owns its lists
    assert all(p["month"][0] != -1 for i, p in enumerate(p_list) if i != 3)


def test_loader_arity_shim():
    """The src/ drivers unpack 5 values from loaders that return 6 in this fork (src/GC/Verify-GC.py:23 vs
    utils/verif_utils.py:241); the drop-in's loaders hand each caller the tuple length it unpacks."""
    from utils.verif_utils import _tolerant_arity

    @_tolerant_arity
    def loader():
        return ("df", 1, 2, 3, 4, "encoders")

    df, a, b, c, d = loader()
    assert (df, d) == ("df", 4)
    df, a, b, c, d, enc = loader()
    assert enc == "encoders"
    whole = loader()
    assert len(whole) == 6

    def five():
        x0, x1, x2, x3, x4 = loader()
        return x4
    assert five() == 4


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["GC-1", "GC-5", "AC-1", "AC-7", "BM-4", "BM-6"])
def test_heuristic_prune_and_mask_helpers(name):
    from utils.prune import compression_ratio, dead_node_from_bound, heuristic_prune, merge_dead_nodes, prune_neurons
    w, b = G.wb(name)
    dims = [w[0].shape[0]] + [x.shape[1] for x in w]
    for c in G.cases(name):
        bounds = tuple(_lists(dims, c[k]) for k in ("ws_lb", "ws_ub", "pl_lb", "pl_ub"))
        cand = _lists(dims, c["cand"], int); b_cand = _lists(dims, c["b_cand"], int); b_dead = _lists(dims, c["b_dead"], int)
        bd, bc, _ = dead_node_from_bound(cand, w, b, None, bounds[1])
        for l in bd:
            if 0 not in l:
                l[0] = 0
        assert bd == b_dead and bc == b_cand
        assert compression_ratio(b_dead) == float(c["b_compression"])
        for pct in (5, 20, 50, 100):
            new, merged = heuristic_prune(bounds, cand, b_cand, merge_dead_nodes(b_dead, [[0] * len(l) for l in b_dead]),
                                          None, pct, w, b)
            assert new == _lists(dims, c[f"h{pct}_new"], int) and merged == _lists(dims, c[f"h{pct}_merged"], int)
        pr_w, pr_b = prune_neurons(w, b, b_dead)
        for l in range(len(w)):
            alive = len(b_dead[l]) - sum(b_dead[l])
            assert pr_w[l].shape[1] == alive and pr_b[l].shape == (alive,)
            if l:
                assert pr_w[l].shape[0] == pr_w[l - 1].shape[1]


def test_model_function_modules_resolve():
    for name in ("AC-1", "GC-5", "BM-13", "AC-16"):
        mod = importlib.import_module(f"utils.{name}-Model-Functions")
        assert callable(mod.layer_net) and callable(mod.net) and callable(mod.z3_net)


def test_star_import_surface():
    """names the drivers resolve through `from utils.x import *`."""
    ns = {}
    exec("from utils.input_partition import *\nfrom utils.verif_utils import *\nfrom utils.prune import *", ns)
    for k in ("partition", "partitioned_ranges", "partition_df", "partitioned_ranges_df", "sound_prune", "sound_prune_bank",
              "sound_prune_german", "sound_prune_compass", "sound_prune_default", "heuristic_prune", "prune_neurons",
              "compression_ratio", "get_y_pred", "sigmoid", "candidate_dead_nodes", "neuron_bounds", "dead_node_from_bound",
              "simluate_data", "merge_dead_nodes", "np", "pd", "os", "time", "copy", "load_model", "accuracy_score", "f1_score",
              "LabelEncoder", "KBinsDiscretizer"):
        assert k in ns, k


# ---------------------------------------------------------------------------------------------------
# GPU
# ---------------------------------------------------------------------------------------------------
SOUND = {"AC": "sound_prune", "GC": "sound_prune_german", "BM": "sound_prune_bank", "CP": "sound_prune_compass"}


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["GC-1", "GC-3", "GC-4", "GC-5", "AC-1", "AC-4", "AC-8", "AC-12", "BM-1", "BM-4", "BM-6", "CP-1"])
def test_sound_prune_returns_what_the_reference_returns(name):
    """seed -> sound_prune*(df, w, b, S, layer_net, p): every element of the 8-tuple against the reference's."""
    import utils.prune as P
    mod = importlib.import_module(f"utils.{name}-Model-Functions")
    w, b = G.wb(name)
    dom = G.domain(name)
    dims = [w[0].shape[0]] + [x.shape[1] for x in w]
    fn = getattr(P, SOUND[dom.name])
    for c in G.cases(name):
        np.random.seed(int(c["seed"]))
        bounds, cand, s_cand, b_dead, s_dead, dead, pos_prob, sim_df = fn(_df(dom), w, b, int(c["S"]), mod.layer_net,
                                                                          _box_dict(dom, c["box"]))
        assert hashlib.sha256(sim_df.to_numpy().astype(np.int32).tobytes()).digest() == c["sim_sha256"].tobytes()
        assert list(sim_df.columns) == dom.columns
        for got, key in zip(bounds, ("ws_lb", "ws_ub", "pl_lb", "pl_ub")):
            assert got == _lists(dims, c[key]), f"{c['tag']} {key}"
            assert type(got[0][0]) is float
        assert cand == _lists(dims, c["cand"], int) and b_dead == _lists(dims, c["b_dead"], int)
        assert pos_prob == _lists(dims, c["pos_prob"])
        if "dead" in c:
            assert s_dead == _lists(dims, c["s_dead"], int) and dead == _lists(dims, c["dead"], int)
            assert s_cand == _lists(dims, c["s_cand"], int)
        assert P.compression_ratio(b_dead) == float(c["b_compression"])


@pytest.mark.gpu
@pytest.mark.parametrize("name", G.singular_models)
def test_sound_prune_with_singular_verification_fixtures(name):
    """Default mode against the unmodified reference's full sound_prune* (Z3 included) on every model with an
    equal-width hidden layer: s_dead, s_candidates, dead_nodes and the compression ratios the drivers print."""
    import utils.prune as P
    mod = importlib.import_module(f"utils.{name}-Model-Functions")
    w, b = G.wb(name)
    dom = G.domain(name)
    dims = [w[0].shape[0]] + [x.shape[1] for x in w]
    fn = getattr(P, SOUND[dom.name])
    for c in G.singular_cases(name):
        np.random.seed(int(c["index"]))
        bounds, cand, s_cand, b_dead, s_dead, dead, pos_prob, sim_df = fn(_df(dom), w, b, int(c["S"]), mod.layer_net,
                                                                          _box_dict(dom, c["box"]))
        assert hashlib.sha256(sim_df.to_numpy().astype(np.int32).tobytes()).digest() == c["sim_sha256"].tobytes()
        assert cand == _lists(dims, c["cand"], int) and b_dead == _lists(dims, c["b_dead"], int), c["tag"]
        assert s_dead == _lists(dims, c["s_dead"], int) and s_cand == _lists(dims, c["s_cand"], int), c["tag"]
        assert dead == _lists(dims, c["dead"], int), c["tag"]
        assert bounds[1] == _lists(dims, c["ws_ub"]) and pos_prob == _lists(dims, c["pos_prob"])
        assert P.compression_ratio(s_dead) == float(c["s_compression"])
        assert P.compression_ratio(dead) == float(c["st_compression"])


@pytest.mark.gpu
def test_sliver_goes_to_the_host_z3_query(monkeypatch):
    """tests/golden/sliver_case.npz (the reference with Z3 on a constructed network whose ws_ub[0][0] = 2^-50): the
    device flags the neuron, the drop-in puts the reference's own query to Z3 for it -- and only for it -- and the
    8-tuple equals the reference's."""
    pytest.importorskip("z3")
    import utils.prune as P
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "sliver_case.npz"))
    w = [z[f"w{l}"] for l in range(3)]; b = [z[f"b{l}"] for l in range(3)]
    dom = G.domain("AC-8")
    dims = [13, 5, 5, 1]
    asked = []
    real = P._z3_positive_is_unsat
    monkeypatch.setattr(P, "_z3_positive_is_unsat", lambda l, j, *a: asked.append((l, j)) or real(l, j, *a))
    mod = importlib.import_module("utils.AC-8-Model-Functions")
    np.random.seed(int(z["seed"]))
    out = P.sound_prune(_df(dom), w, b, int(z["S"]), mod.layer_net, _box_dict(dom, z["box"]))
    assert asked == [(0, 0)]
    assert hashlib.sha256(out[7].to_numpy().astype(np.int32).tobytes()).digest() == z["sim_sha256"].tobytes()
    for got, key in zip(out[0], ("ws_lb", "ws_ub", "pl_lb", "pl_ub")):
        assert got == _lists(dims, z[key]), key
    for got, key in ((out[1], "cand"), (out[2], "s_cand"), (out[3], "b_dead"), (out[4], "s_dead"), (out[5], "dead")):
        assert got == _lists(dims, z[key], int), key
    assert out[6] == _lists(dims, z["pos_prob"])


@pytest.mark.gpu
def test_window_batching_is_invisible(monkeypatch):
    """sound_prune called partition by partition over a shuffled PartitionList: with the window on (one launch per 16
    partitions here) and off, the 8-tuples, the samples and the state of numpy's global stream after every call are
    identical -- also when the caller draws random numbers of its own between two calls and when it abandons a window
    halfway (new model)."""
    import random
    import utils.prune as P
    from utils.input_partition import partition, partitioned_ranges
    dom = G.domain("AC-1")
    mod = importlib.import_module("utils.AC-1-Model-Functions")
    runs = {}
    for window in ("16", "1"):
        monkeypatch.setenv("FAIRIFY_GPU_WINDOW", window)
        P._window = None
        rd = dom.copy_ranges()
        p_list = partitioned_ranges(rd.keys(), dom.pa, partition(rd, dom.threshold), rd)
        random.Random(7).shuffle(p_list)
        np.random.seed(123)
        trace = []
        for name, count in (("AC-1", 40), ("AC-10", 21), ("AC-1", 5)):
            w, b = G.wb(name)
            for k, p in enumerate(p_list[:count]):
                out = P.sound_prune(_df(dom), w, b, 200, mod.layer_net, p)
                if k % 7 == 3:
                    np.random.randint(0, 10, size=3)                  # the caller's own draws
                trace.append((out[0], out[1], out[2], out[3], out[4], out[5], out[6], out[7].to_numpy().tobytes(),
                              np.random.get_state()[1].tobytes(), np.random.get_state()[2]))
        runs[window] = trace
    assert len(runs["16"]) == len(runs["1"]) == 66
    for k, (a, b_) in enumerate(zip(runs["16"], runs["1"])):
        assert a == b_, f"call {k} differs between the batched and the one-at-a-time path"


@pytest.mark.gpu
def test_driver_partition_loop_runs_on_the_dropin():
    """The body of the drivers' partition loop (src/GC/Verify-GC.py:106-320, same call sequence and the same names the
    drivers star-import): sound_prune_german -> compression_ratio x3 -> prune_neurons -> z3_net on the pruned weights
    -> Z3 check of the fairness property -> heuristic_prune on unknown -> C-check with net -> get_y_pred x3.  The
    reference checkout (and with it the unmodified script) is not on the GPU box, so the loop is restated here; Z3 runs
    on the host exactly as in the driver."""
    z3 = pytest.importorskip("z3")
    ns = {}
    exec("from utils.input_partition import *\nfrom utils.verif_utils import *\nfrom utils.prune import *", ns)
    mod = importlib.import_module("utils.GC-4-Model-Functions")
    dom = G.domain("GC-4")
    w, b = G.wb("GC-4")
    df = _df(dom)
    range_dict = dom.copy_ranges()
    A = range_dict.keys(); PA = dom.pa
    p_list = ns["partitioned_ranges"](A, PA, ns["partition"](range_dict, dom.threshold), range_dict)
    assert len(p_list) == 201
    import random
    random.Random(3).shuffle(p_list)
    np.random.seed(11)
    verdicts = []
    for p in p_list[:6]:
        bounds, candidates, s_candidates, b_deads, s_deads, st_deads, pos_prob, sim_X_df = \
            ns["sound_prune_german"](df, w, b, 1000, mod.layer_net, p)
        ratios = [ns["compression_ratio"](x) for x in (b_deads, s_deads, st_deads)]
        assert 0.0 <= ratios[0] <= ratios[2] <= 1.0
        pr_w, pr_b = ns["prune_neurons"](w, b, st_deads)
        x = np.array([z3.Int('x%s' % i) for i in range(20)]); x_ = np.array([z3.Int('x_%s' % i) for i in range(20)])
        y = mod.z3_net(x, pr_w, pr_b); y_ = mod.z3_net(x_, pr_w, pr_b)
        s = z3.Solver(); s.set("timeout", 20000)
        for i, col in enumerate(dom.columns):
            lo, hi = p[col]
            s.add(x[i] >= lo, x[i] <= hi, x_[i] >= lo, x_[i] <= hi)
            if col not in PA:
                s.add(x[i] == x_[i])
            else:
                s.add(x[i] != x_[i])
        s.add(z3.Or(z3.And(y[0] < 0, y_[0] > 0), z3.And(y[0] > 0, y_[0] < 0)))
        res = s.check()
        if res == z3.unknown:
            h_deads, deads = ns["heuristic_prune"](bounds, candidates, s_candidates, st_deads, pos_prob, 5, w, b)
            pr_w, pr_b = ns["prune_neurons"](w, b, deads)
        if res == z3.sat:
            mdl = s.model()
            d1 = np.array([mdl.eval(v, model_completion=True).as_long() for v in x], dtype=np.float32)
            d2 = np.array([mdl.eval(v, model_completion=True).as_long() for v in x_], dtype=np.float32)
            # C-check (src/GC/Verify-GC.py:229-249): the counterexample flips the class on the pruned and the original net
            c1 = ns["sigmoid"](mod.net(d1, pr_w, pr_b)) > 0.5; c2 = ns["sigmoid"](mod.net(d2, pr_w, pr_b)) > 0.5
            o1 = ns["sigmoid"](mod.net(d1, w, b)) > 0.5; o2 = ns["sigmoid"](mod.net(d2, w, b)) > 0.5
            assert c1 != c2 and o1 != o2
        sim_X = sim_X_df.to_numpy()
        sim_y_orig = ns["get_y_pred"](mod.net, w, b, sim_X)
        sim_y = ns["get_y_pred"](mod.net, pr_w, pr_b, sim_X)
        if res != z3.unknown:
            assert ns["accuracy_score"](sim_y, sim_y_orig) == 1.0          # 'Pruned-acc' of sound pruning
        verdicts.append(str(res))
    assert len(verdicts) == 6 and set(verdicts) <= {"sat", "unsat", "unknown"}


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["GC-3", "AC-1"])
def test_reference_mode_runs_the_z3_queries(name, monkeypatch):
    """FAIRIFY_GPU_SINGULAR=reference issues the reference's per-neuron Z3 query for every remaining candidate."""
    pytest.importorskip("z3")
    import utils.prune as P
    monkeypatch.setenv("FAIRIFY_GPU_SINGULAR", "reference")
    mod = importlib.import_module(f"utils.{name}-Model-Functions")
    w, b = G.wb(name); dom = G.domain(name)
    dims = [w[0].shape[0]] + [x.shape[1] for x in w]
    c = next(iter(G.cases(name)))
    np.random.seed(int(c["seed"]))
    out = getattr(P, SOUND[dom.name])(_df(dom), w, b, int(c["S"]), mod.layer_net, _box_dict(dom, c["box"]))
    assert out[4] == _lists(dims, c["s_dead"], int) and out[5] == _lists(dims, c["dead"], int)
    assert out[2] == _lists(dims, c["s_cand"], int)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["GC-2", "AC-3", "BM-8"])
def test_pieces_and_predictions(name):
    import utils.prune as P
    from utils.verif_utils import get_y_pred
    mod = importlib.import_module(f"utils.{name}-Model-Functions")
    w, b = G.wb(name); dom = G.domain(name)
    dims = [w[0].shape[0]] + [x.shape[1] for x in w]
    c = next(iter(G.cases(name)))
    x_df = _df(dom).drop(labels=[dom.label], axis=1)
    np.random.seed(int(c["seed"]))
    sim_df = P.simluate_data(x_df, int(c["S"]), _box_dict(dom, c["box"]))
    cand, prob = P.candidate_dead_nodes(sim_df.to_numpy(), w, b, mod.layer_net)
    assert cand == _lists(dims, c["cand"], int) and prob == _lists(dims, c["pos_prob"])
    nb = P.neuron_bounds(sim_df, w, b, _box_dict(dom, c["box"]))
    assert nb[1] == _lists(dims, c["ws_ub"]) and nb[2] == _lists(dims, c["pl_lb"])
    rows = sim_df.to_numpy()[:64]
    pred = get_y_pred(mod.net, w, b, rows)
    assert [bool(p[0]) for p in pred] == list(c["logit64"] > 0)
    # single-row calls keep the reference's shapes and dtypes
    y = mod.net(rows[0], w, b)
    assert y.shape == (1,) and y.dtype == np.float64 and (y[0] > 0) == (c["logit64"][0] > 0)
    y32 = mod.net(rows[0].astype(np.float32), w, b)
    assert y32.dtype == np.float32 and (y32[0] > 0) == (c["logit32"][0] > 0)
    layers = mod.layer_net(rows[0], w, b)
    assert [len(l) for l in layers] == dims[1:] and np.isclose(layers[-1][0], c["logit64"][0], rtol=1e-12, atol=1e-12)
    # pruned-vs-original agreement on the partition's own samples ('Pruned-acc' == 1.0, src/AC/Verify-AC.py:269-274)
    pr_w, pr_b = P.prune_neurons(w, b, _lists(dims, c["b_dead"], int))
    a = get_y_pred(mod.net, w, b, sim_df.to_numpy()); p = get_y_pred(mod.net, pr_w, pr_b, sim_df.to_numpy())
    assert all(bool(x[0]) == bool(y_[0]) for x, y_ in zip(a, p))
