"""The drop-in `utils` package (same names / signatures / return structures as the reference's utils/*):
host-only pieces on CPU, the GPU-backed entry points under the gpu marker."""
import hashlib
import importlib

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
        else:
            assert dead == _lists(dims, c["b_dead"], int) and s_cand == _lists(dims, c["b_cand"], int)
        assert P.compression_ratio(b_dead) == float(c["b_compression"])


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
