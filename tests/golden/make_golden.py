#!/usr/bin/env python3
"""Generate the committed golden fixtures by running the UNMODIFIED reference in the build container.

    python tests/golden/make_golden.py            # writes tests/golden/*.npz  (needs /root/reference)

What it runs (reference file:line):
  * utils/input_partition.py:17-76  partition + partitioned_ranges  -> grid checksums / sampled boxes
  * utils/prune.py:205-222          simluate_data (after np.random.seed(k))
  * utils/prune.py:168-192          candidate_dead_nodes with utils/<M>-Model-Functions.py layer_net
  * utils/prune.py:105-164          neuron_bounds   (weights pre-cast to float64: the float64 semantics of
                                    the reference's pinned numpy 1.19; numpy >= 2 would otherwise keep
                                    float32 products)
  * utils/prune.py:226-251          dead_node_from_bound, :941-948 merge_dead_nodes, :194-203 compression_ratio
  * utils/prune.py:671-783          sound_prune / sound_prune_german / sound_prune_bank INCLUDING the Z3
                                    singular_verification, on a few small models
  * utils/prune.py:862-939          heuristic_prune
  * utils/<M>-Model-Functions.py    net (float64 on integer rows, float32 on float32 rows)
  * src/AC/metrics.py:40-264        CausalDiscriminationDetector with a numpy float32 predict function
Weights come from models/*/*.h5 through fairify_b200/h5weights.py (no tensorflow / h5py here).
The fixtures hold inputs by seed (the sample matrix is regenerated with the stream-identical
np.random.randint(lo, hi+1, size=(S, d)) and verified here against simluate_data) and outputs by value.
"""
import glob
import hashlib
import os
import sys
import types

import numpy as np
import pandas as pd

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.abspath(os.path.join(HERE, "..", ".."))

# load this repo's pure-data helpers WITHOUT importing its `utils` package
import importlib.util


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


h5weights = _load(os.path.join(ROOT, "fairify_b200", "h5weights.py"), "h5weights")
domains = _load(os.path.join(ROOT, "fairify_b200", "domains.py"), "domains")
refshim = _load(os.path.join(ROOT, "oracle", "refshim.py"), "refshim")

REF = refshim.REFERENCE_ROOT
MODEL_DIRS = {"AC": "adult", "GC": "german", "BM": "bank", "CP": "compass", "DF": "default"}
# second, finer partition threshold per domain for the prune cases (grid sizes 67 392 / 2 001 / 6 300 / 4)
FINER_T = {"AC": 8, "GC": 30, "BM": 40, "CP": 10}


def flat(lists):
    return np.concatenate([np.asarray(l, dtype=np.float64).ravel() for l in lists])


def flat_u8(lists):
    return np.concatenate([np.asarray(l, dtype=np.uint8).ravel() for l in lists])


def box_array(p, columns):
    return np.array([[p[c][0], p[c][1]] for c in columns], dtype=np.int32)


def write_weights():
    out = {}
    for f in sorted(glob.glob(os.path.join(REF, "models", "*", "*.h5"))):
        name = os.path.basename(f)[:-3]
        w, b = h5weights.read_keras_dense_weights(f)
        out[f"{name}/L"] = np.int32(len(w))
        for l, (wl, bl) in enumerate(zip(w, b)):
            out[f"{name}/w{l}"] = wl
            out[f"{name}/b{l}"] = bl
    np.savez_compressed(os.path.join(HERE, "weights.npz"), **out)
    print("weights.npz:", len([k for k in out if k.endswith('/L')]), "models")
    return out


def grids(ref):
    out = {}
    cases = [
        ("AC_T10", "AC", 10, {}), ("GC_T100", "GC", 100, {}), ("BM_T100", "BM", 100, {}),
        ("GC_T10", "GC", 10, {}), ("CP_T5", "CP", 5, {}),
        ("AC_T6_age30_35", "AC", 6, {"age": [30, 35]}),              # targeted/AC/Verify-AC.py:33
        ("GC_T10_purpose7", "GC", 10, {"purpose": [7, 7]}),          # targeted GC
        ("BM_T10_small", "BM", 10, {"duration": [0, 95], "pdays": [0, 40]}),
    ]
    for tag, dom, T, override in cases:
        D = domains.DOMAINS[dom]
        rd = D.copy_ranges()
        for k, v in override.items():
            rd[k] = list(v)
        p_dict = ref.input_partition.partition(rd, T)
        p_list = ref.input_partition.partitioned_ranges(rd.keys(), D.pa, p_dict, rd)
        boxes = np.stack([box_array(p, D.columns) for p in p_list])
        out[f"{tag}/n"] = np.int64(len(p_list))
        out[f"{tag}/sha256"] = np.frombuffer(hashlib.sha256(boxes.tobytes()).digest(), dtype=np.uint8)
        idx = np.unique(np.concatenate([np.arange(min(4, len(p_list))),
                                        np.linspace(0, len(p_list) - 1, 9).astype(np.int64)]))
        out[f"{tag}/idx"] = idx
        out[f"{tag}/boxes"] = boxes[idx]
        out[f"{tag}/T"] = np.int64(T)
        out[f"{tag}/lo"] = np.array([rd[c][0] for c in D.columns], dtype=np.int64)
        out[f"{tag}/hi"] = np.array([rd[c][1] for c in D.columns], dtype=np.int64)
        out[f"{tag}/key_order"] = np.array([D.columns.index(k) for k in rd.keys()], dtype=np.int32)
        print(f"grid {tag}: n={len(p_list)}")
    np.savez_compressed(os.path.join(HERE, "grids.npz"), **out)


def prune_cases(ref, weights):
    P = ref.prune
    out = {}
    models = (["GC-%d" % i for i in range(1, 6)] + ["AC-%d" % i for i in range(1, 13)]
              + ["BM-%d" % i for i in range(1, 9)] + ["CP-1"])
    full_z3 = {"GC-3", "GC-4", "AC-1", "AC-8", "BM-6"}
    sound = {"AC": P.sound_prune, "GC": P.sound_prune_german, "BM": P.sound_prune_bank, "CP": P.sound_prune_compass}
    case_id = 0
    for name in models:
        dom = name.split("-")[0]
        D = domains.DOMAINS[dom]
        L = int(weights[f"{name}/L"])
        w32 = [weights[f"{name}/w{l}"] for l in range(L)]
        b32 = [weights[f"{name}/b{l}"] for l in range(L)]
        w64 = [x.astype(np.float64) for x in w32]
        b64 = [x.astype(np.float64) for x in b32]
        layer_net, net = refshim.model_functions(name)
        df = pd.DataFrame(np.zeros((2, D.d + 1)), columns=D.columns + [D.label])
        x_df = df.drop(labels=[D.label], axis=1)
        rd = D.copy_ranges()
        # two grids per domain: the src/ threshold and a finer one
        plist = []
        for T in (D.threshold, FINER_T[dom]):
            p_dict = ref.input_partition.partition(rd, T)
            pl = ref.input_partition.partitioned_ranges(rd.keys(), D.pa, p_dict, rd)
            sel = sorted(set([0, 1, len(pl) // 3, len(pl) // 2, len(pl) - 1]))
            plist += [pl[i] for i in sel]
        big = sum(x.size for x in w32) > 4000
        if big:
            plist = plist[:4]
        for k, p in enumerate(plist):
            S = 1000 if k % 3 == 0 else 200
            seed = 1000 + case_id
            tag = f"{name}/{k}"
            np.random.seed(seed)
            with refshim.quiet():
                sim_df = P.simluate_data(x_df, S, p)
            sim = sim_df.to_numpy()
            # the vectorised draw the tests use must be stream-identical to the scalar loop
            np.random.seed(seed)
            lo = np.array([p[c][0] for c in D.columns]); hi = np.array([p[c][1] for c in D.columns])
            vec = np.random.randint(lo, hi + 1, size=(S, D.d))
            assert np.array_equal(vec, sim), "vectorised randint is not stream-identical"
            with refshim.quiet():
                cand, pos_prob = P.candidate_dead_nodes(sim, w64, b64, layer_net)
                ws_lb, ws_ub, pl_lb, pl_ub = P.neuron_bounds(sim_df, w64, b64, p)
                b_dead, b_cand, b_ratio = P.dead_node_from_bound(cand, w64, b64, p, ws_ub)
            for l in b_dead:
                if 0 not in l:
                    l[0] = 0
            assert isinstance(ws_ub[0][0], (float, np.float64))
            out[f"{tag}/seed"] = np.int64(seed); out[f"{tag}/S"] = np.int64(S)
            out[f"{tag}/box"] = box_array(p, D.columns)
            out[f"{tag}/sim_sha256"] = np.frombuffer(hashlib.sha256(sim.astype(np.int32).tobytes()).digest(), dtype=np.uint8)
            out[f"{tag}/cand"] = flat_u8(cand); out[f"{tag}/pos_prob"] = flat(pos_prob)
            out[f"{tag}/ws_lb"] = flat(ws_lb); out[f"{tag}/ws_ub"] = flat(ws_ub)
            out[f"{tag}/pl_lb"] = flat(pl_lb); out[f"{tag}/pl_ub"] = flat(pl_ub)
            out[f"{tag}/b_dead"] = flat_u8(b_dead); out[f"{tag}/b_cand"] = flat_u8(b_cand)
            out[f"{tag}/b_compression"] = np.float64(P.compression_ratio(b_dead))
            # heuristic prune on the bound-only masks (s_candidates == b_candidates when Z3 finds nothing)
            with refshim.quiet():
                merged0 = P.merge_dead_nodes(b_dead, [[0] * len(l) for l in b_dead])
            for pct in (5, 20, 50, 100):
                with refshim.quiet():
                    nd, md = P.heuristic_prune((ws_lb, ws_ub, pl_lb, pl_ub), cand, b_cand, merged0, pos_prob, pct, w64, b64)
                out[f"{tag}/h{pct}_new"] = flat_u8(nd); out[f"{tag}/h{pct}_merged"] = flat_u8(md)
            # logits: float64 on the integer sample rows, float32 on float32 rows with float32 weights
            rows = sim[:64]
            out[f"{tag}/logit64"] = np.array([float(net(r, w64, b64)[0]) for r in rows])
            l32 = [net(r.astype(np.float32), w32, b32)[0] for r in rows]
            assert all(isinstance(v, np.float32) for v in l32)
            out[f"{tag}/logit32"] = np.array(l32, dtype=np.float32)
            if name in full_z3 and k < 3:
                np.random.seed(seed)
                with refshim.quiet():
                    (bounds, c2, s_cand, bd2, s_dead, dead, pp2, sdf2) = sound[dom](df, w64, b64, S, layer_net, p)
                assert np.array_equal(sdf2.to_numpy(), sim) and c2 == cand and bd2 == b_dead
                out[f"{tag}/s_cand"] = flat_u8(s_cand); out[f"{tag}/s_dead"] = flat_u8(s_dead)
                out[f"{tag}/dead"] = flat_u8(dead)
                out[f"{tag}/st_compression"] = np.float64(P.compression_ratio(dead))
                out[f"{tag}/s_compression"] = np.float64(P.compression_ratio(s_dead))
            case_id += 1
        out[f"{name}/ncases"] = np.int64(len(plist))
        print(f"prune cases {name}: {len(plist)}")
    out["models"] = np.array(models)
    np.savez_compressed(os.path.join(HERE, "prune_cases.npz"), **out)


def np_stats():
    """numpy mean / median / percentile on random float64 vectors (heuristic_prune's statistics)."""
    rng = np.random.RandomState(7)
    out = {}
    lens = list(range(1, 41)) + [63, 64, 65, 100, 127, 128, 129, 130, 137, 150, 200, 301]
    vecs, res = [], []
    for n in lens:
        v = rng.standard_normal(n) * 10.0 ** rng.randint(-3, 4)
        vecs.append(v)
        row = [np.mean(v), np.median(v)]
        for pct in (0, 5, 20, 50, 80, 95, 100):
            row.append(np.percentile(v, pct))
        res.append(row)
    out["lens"] = np.array(lens); out["data"] = np.concatenate(vecs); out["res"] = np.array(res)
    out["numpy_version"] = np.array(np.__version__)
    np.savez_compressed(os.path.join(HERE, "np_stats.npz"), **out)


def causal_cases(weights):
    M = refshim.metrics_module("AC")
    out = {}

    def make_predict(name, columns):
        L = int(weights[f"{name}/L"])
        w = [weights[f"{name}/w{l}"] for l in range(L)]
        b = [weights[f"{name}/b{l}"] for l in range(L)]
        stats = {"min_abs": np.inf}

        def predict(fd):
            # the shape of model_predict_fn_original, src/AC/metrics.py:323-325, with the Dense stack in numpy float32
            x = np.array([[fd[f] for f in columns]], dtype=np.float32)
            for l in range(L):
                x = x @ w[l] + b[l]
                if l < L - 1:
                    x = np.maximum(x, np.float32(0))
            stats["min_abs"] = min(stats["min_abs"], float(abs(x[0][0])))
            sig = np.float32(1) / (np.float32(1) + np.exp(-x[0][0], dtype=np.float32))
            return int(sig > 0.5)
        return predict, stats

    cases = [("AC-1", "AC", ["sex"], 0.999, 0.0001), ("AC-3", "AC", ["race"], 0.99, 0.01),
             ("GC-1", "GC", ["age"], 0.999, 0.0001), ("BM-1", "BM", ["age"], 0.99, 0.02),
             ("AC-1", "AC", ["race", "sex"], 0.95, 0.02), ("GC-4", "GC", ["sex", "age"], 0.999, 0.0001)]
    for ci, (name, dom, pa, conf, margin) in enumerate(cases):
        D = domains.DOMAINS[dom]
        predict, stats = make_predict(name, D.columns)
        det = M.CausalDiscriminationDetector(predict, max_samples=1000, min_samples=100)
        for c in D.columns:
            lo, hi = D.ranges[c]
            vals = list(range(lo, hi + 1))
            if len(vals) > 60:                       # keep value lists dataset-like
                vals = vals[::max(1, len(vals) // 40)]
            det.add_feature(c, vals)
        test_cases, rate, pairs = det.causal_discrimination(pa, conf=conf, margin=margin)
        tag = f"case{ci}"
        out[f"{tag}/model"] = np.array(name); out[f"{tag}/pa"] = np.array(pa)
        out[f"{tag}/conf"] = np.float64(conf); out[f"{tag}/margin"] = np.float64(margin)
        out[f"{tag}/rate"] = np.float64(rate)
        out[f"{tag}/test_cases"] = np.array([[float(tc[c]) for c in D.columns] for tc in test_cases], dtype=np.float32)
        out[f"{tag}/pairs"] = np.array([[[float(a[c]) for c in D.columns], [float(b_[c]) for c in D.columns]]
                                        for a, b_ in pairs], dtype=np.float32).reshape(-1, 2, D.d)
        out[f"{tag}/min_abs_logit"] = np.float64(stats["min_abs"])
        for c in D.columns:
            out[f"{tag}/values/{c}"] = np.array([float(v) for v in det.inputs[c].values], dtype=np.float32)
        print(f"causal {name} {pa}: rate={rate:.4f} cases={len(test_cases)} pairs={len(pairs)} min|logit|={stats['min_abs']:.3g}")
    out["ncases"] = np.int64(len(cases))
    np.savez_compressed(os.path.join(HERE, "causal_cases.npz"), **out)


def main():
    if not refshim.available():
        sys.exit("reference checkout not found; fixtures can only be regenerated in the build container")
    ref = refshim.load()
    weights = write_weights()
    grids(ref)
    np_stats()
    prune_cases(ref, weights)
    causal_cases(weights)
    density_case(ref)



def density_case(ref):
    """p_list_density (utils/input_partition.py:184-218) on a synthetic Adult frame: rows per partition / rows."""
    D = domains.DOMAINS["AC"]
    rng = np.random.RandomState(11)
    out = {}
    for tag, T, nrows in (("AC_T30", 30, 400), ("AC_T12", 12, 250)):
        rd = D.copy_ranges()
        p_dict = ref.input_partition.partition(rd, T)
        p_list = ref.input_partition.partitioned_ranges(rd.keys(), D.pa, p_dict, rd)
        lo = np.array([rd[c][0] for c in D.columns]); hi = np.array([rd[c][1] for c in D.columns])
        X = rng.randint(lo, hi + 1, size=(nrows, D.d))
        X[::17, 0] = 5                      # a few rows below the age range: they fall in no partition
        X[5::23, 11] = 150                  # and above the hours-per-week range
        df = pd.DataFrame(np.column_stack([X, rng.randint(0, 2, size=nrows)]), columns=D.columns + [D.label])
        dens = ref.input_partition.p_list_density(D.copy_ranges(), p_list, df)     # (the reference overwrites its range_dict argument)
        out[f"{tag}/T"] = np.int64(T); out[f"{tag}/X"] = X.astype(np.int64); out[f"{tag}/density"] = np.array(dens)
        out[f"{tag}/n"] = np.int64(len(p_list))
        print(f"density {tag}: {len(p_list)} partitions, {nrows} rows, covered {sum(dens):.3f}")
    np.savez_compressed(os.path.join(HERE, "density.npz"), **out)


# ----------------------------------------------------------------------------------------------------
# full sound_prune* INCLUDING the Z3 singular_verification (utils/prune.py:276-644) on every model with a
# hidden layer whose input and output widths are equal.  There the reference's query bounds the inputs of layer
# l >= 1 with that layer's OWN post-activation bounds (intermediate_domain_constraint is called with
# layer_index, :266-273, :302-345), so it answers unsat -- s_dead = 1 -- for neurons whose interval upper bound is
# far above 0.  Z3 is slow (0.05-0.5 s per neuron), so the cases are spread over worker processes.
#     python tests/golden/make_golden.py singular        # writes tests/golden/singular_cases.npz
# ----------------------------------------------------------------------------------------------------
# (CP-2..CP-10 take 12 features, but singular_verification_compass hard-codes 6 -- z3_layer1_ws_net_compass,
#  utils/prune.py:74-79 -- and raises IndexError on them; the 32-32 COMPAS shape is covered by CP-11, d = 6.)
SINGULAR_MODELS = ["AC-4", "AC-5", "AC-6", "AC-8", "AC-9", "AC-10", "AC-11", "AC-12", "BM-6", "BM-7", "CP-11",
                   "AC-1", "GC-4", "BM-8"]          # the last three: funnel layers (inputs k >= out stay unbounded)
SINGULAR_PER_MODEL = 20
SINGULAR_FORCED = {"AC-10": [13529, 5595]}            # partitions on which the verdict of round 1 saw s_dead = 1
_sing_state = {}


def _singular_domain(name):
    return domains.DOMAINS[name.split("-")[0]]


def _singular_worker(task):
    name, idx, S = task
    if "ref" not in _sing_state:
        _sing_state["ref"] = refshim.load()
        _sing_state["weights"] = np.load(os.path.join(HERE, "weights.npz"))
    ref, weights = _sing_state["ref"], _sing_state["weights"]
    P = ref.prune
    dom = name.split("-")[0]
    sound = {"AC": P.sound_prune, "GC": P.sound_prune_german, "BM": P.sound_prune_bank, "CP": P.sound_prune_compass}[dom]
    D = _singular_domain(name)
    L = int(weights[f"{name}/L"])
    w64 = [weights[f"{name}/w{l}"].astype(np.float64) for l in range(L)]
    b64 = [weights[f"{name}/b{l}"].astype(np.float64) for l in range(L)]
    layer_net, _ = refshim.model_functions(name)
    df = pd.DataFrame(np.zeros((2, D.d + 1)), columns=D.columns + [D.label])
    rd = D.copy_ranges()
    key = (name, "plist")
    if key not in _sing_state:
        p_dict = ref.input_partition.partition(rd, D.threshold)
        _sing_state[key] = ref.input_partition.partitioned_ranges(rd.keys(), D.pa, p_dict, rd)
    p = _sing_state[key][idx]
    np.random.seed(idx)
    with refshim.quiet():
        bounds, cand, s_cand, b_dead, s_dead, dead, pos_prob, sim_df = sound(df, w64, b64, S, layer_net, p)
    ws_lb, ws_ub, pl_lb, pl_ub = bounds
    sim = sim_df.to_numpy()
    return {"model": name, "index": idx, "S": S, "box": box_array(p, D.columns),
            "sim_sha256": np.frombuffer(hashlib.sha256(sim.astype(np.int32).tobytes()).digest(), dtype=np.uint8),
            "cand": flat_u8(cand), "s_cand": flat_u8(s_cand), "b_dead": flat_u8(b_dead), "s_dead": flat_u8(s_dead),
            "dead": flat_u8(dead), "ws_ub": flat(ws_ub), "pl_lb": flat(pl_lb), "pl_ub": flat(pl_ub),
            "pos_prob": flat(pos_prob), "s_compression": np.float64(P.compression_ratio(s_dead)),
            "st_compression": np.float64(P.compression_ratio(dead))}


def singular_cases():
    import multiprocessing as mp
    weights = np.load(os.path.join(HERE, "weights.npz"))
    tasks = []
    for name in SINGULAR_MODELS:
        D = _singular_domain(name)
        rd = D.copy_ranges()
        n = 1
        for lo, hi in rd.values():
            size = hi - lo + 1
            n *= 1 if size <= D.threshold else -(-size // D.threshold)
        rng = np.random.RandomState(sum(map(ord, name)))
        count = 40 if name == "AC-5" else SINGULAR_PER_MODEL
        idxs = list(SINGULAR_FORCED.get(name, []))
        while len(idxs) < min(count, n):
            k = int(rng.randint(0, n))
            if k not in idxs:
                idxs.append(k)
        tasks += [(name, k, 200) for k in idxs]
    # biggest models first so the pool drains evenly
    order = {m: i for i, m in enumerate(["AC-4", "BM-7", "AC-5", "CP-11", "AC-11", "AC-12"])}
    tasks.sort(key=lambda t: order.get(t[0], 99))
    with mp.get_context("fork").Pool(min(8, os.cpu_count() or 1)) as pool:
        results = pool.map(_singular_worker, tasks, chunksize=1)
    out = {}
    per_model = {}
    for r in results:
        k = per_model.setdefault(r["model"], 0)
        per_model[r["model"]] = k + 1
        tag = f"{r['model']}/{k}"
        for key, v in r.items():
            if key != "model":
                out[f"{tag}/{key}"] = np.asarray(v)
    for m, k in per_model.items():
        out[f"{m}/ncases"] = np.int64(k)
        hits = sum(int(out[f"{m}/{i}/s_dead"].sum()) for i in range(k))
        print(f"singular cases {m}: {k} partitions, {hits} neurons with s_dead = 1")
    out["models"] = np.array(list(per_model.keys()))
    np.savez_compressed(os.path.join(HERE, "singular_cases.npz"), **out)


def sliver_network():
    """A 13-5-5-1 network (AC-8's shape, Adult columns) whose layer-0 neuron 0 has ws_ub = 2^-50 on the box below: the
    terms 2^-52 * 4 + 1 * 4 - 4 cancel to within the rounding-error tolerance, so the GPU path must flag the neuron
    and hand it to Z3."""
    w0 = np.zeros((13, 5), np.float32); w0[0, 0] = 2.0 ** -52; w0[1, 0] = 1.0; w0[0, 1:] = 1.0; w0[1, 1:] = 0.5
    b0 = np.array([-4.0, 0.0, 1.0, -1.0, 0.5], np.float32)
    rng = np.random.RandomState(5)
    w1 = (rng.standard_normal((5, 5)) * 0.5).astype(np.float32); b1 = (rng.standard_normal(5) * 0.5).astype(np.float32)
    w2 = rng.standard_normal((5, 1)).astype(np.float32); b2 = np.array([0.1], np.float32)
    return [w0, w1, w2], [b0, b1, b2]


def sliver_case():
    """The reference's sound_prune (Z3 included) on sliver_network(): python tests/golden/make_golden.py sliver"""
    ref = refshim.load()
    P = ref.prune
    D = domains.DOMAINS["AC"]
    w32, b32 = sliver_network()
    w64 = [x.astype(np.float64) for x in w32]; b64 = [x.astype(np.float64) for x in b32]
    layer_net, _ = refshim.model_functions("AC-8")
    df = pd.DataFrame(np.zeros((2, D.d + 1)), columns=D.columns + [D.label])
    p = D.copy_ranges()
    p["age"] = [0, 4]; p["workclass"] = [0, 4]
    out = {}
    S = 6
    for seed in range(1000):                      # a seed whose samples never activate neuron 0 (needs x1 = 4, x0 >= 1)
        np.random.seed(seed)
        with refshim.quiet():
            bounds, cand, s_cand, b_dead, s_dead, dead, pos_prob, sim_df = P.sound_prune(df, w64, b64, S, layer_net, p)
        if cand[0][0] == 1:
            break
    assert cand[0][0] == 1 and bounds[1][0][0] == 2.0 ** -50 and b_dead[0][0] == 0
    out["seed"] = np.int64(seed); out["S"] = np.int64(S); out["box"] = box_array(p, D.columns)
    sim = sim_df.to_numpy()
    out["sim_sha256"] = np.frombuffer(hashlib.sha256(sim.astype(np.int32).tobytes()).digest(), dtype=np.uint8)
    for l in range(3):
        out[f"w{l}"] = w32[l]; out[f"b{l}"] = b32[l]
    for k, v in (("cand", cand), ("s_cand", s_cand), ("b_dead", b_dead), ("s_dead", s_dead), ("dead", dead)):
        out[k] = flat_u8(v)
    for k, v in zip(("ws_lb", "ws_ub", "pl_lb", "pl_ub"), bounds):
        out[k] = flat(v)
    out["pos_prob"] = flat(pos_prob)
    np.savez_compressed(os.path.join(HERE, "sliver_case.npz"), **out)
    print("sliver case: seed", seed, "cand", flat_u8(cand), "s_dead", flat_u8(s_dead), "dead", flat_u8(dead))


def hybrid_case():
    """hybrid_predict / find_partition_result_for_point of src/AC/Verify-AC-experiment-new2.py:587-638, taken out of
    that script at generation time (it is a module-level program, not importable) and run on a random verdict table:
    python tests/golden/make_golden.py hybrid  ->  tests/golden/hybrid_case.npz"""
    import ast
    path = os.path.join(REF, "src", "AC", "Verify-AC-experiment-new2.py")
    tree = ast.parse(open(path).read())
    wanted = ("partition_to_key", "key_to_partition", "find_partition_result_for_point", "hybrid_predict")
    ns = {"np": np}
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in wanted:
            exec(compile(ast.Module([node], []), path, "exec"), ns)
    assert all(k in ns for k in wanted)
    ref = refshim.load()
    weights = np.load(os.path.join(HERE, "weights.npz"))
    D = domains.DOMAINS["AC"]

    class Net:          # what hybrid_predict uses of a keras model: predict(x[1, d]) -> sigmoid outputs, float32 Dense stack
        def __init__(self, name):
            L = int(weights[f"{name}/L"])
            self.w = [weights[f"{name}/w{l}"] for l in range(L)]; self.b = [weights[f"{name}/b{l}"] for l in range(L)]
            self.logits = []

        def predict(self, x, verbose=0):
            x = np.asarray(x, dtype=np.float32)
            for l in range(len(self.w)):
                x = x @ self.w[l] + self.b[l]
                if l < len(self.w) - 1:
                    x = np.maximum(x, np.float32(0))
            self.last_logit = np.float32(x[0][0])
            return np.float32(1) / (np.float32(1) + np.exp(-x, dtype=np.float32))

    orig, fair = Net("AC-3"), Net("AC-1")
    rd = D.copy_ranges()
    T = 10
    p_dict = ref.input_partition.partition(rd, T)
    p_list = ref.input_partition.partitioned_ranges(rd.keys(), D.pa, p_dict, rd)
    rng = np.random.RandomState(21)
    status = np.zeros(len(p_list), dtype=np.uint8)
    names = {1: "sat", 2: "unsat", 3: "unknown"}
    partition_results = {}
    lo = np.array([rd[c][0] for c in D.columns]); hi = np.array([rd[c][1] for c in D.columns])
    X = rng.randint(lo, hi + 1, size=(300, D.d)).astype(np.float64)
    X[::29, 0] = 5.0                                   # below the age range: in no partition
    # verdicts for the partitions the points fall in (two thirds of them) and for some others
    boxes = np.stack([box_array(p, D.columns) for p in p_list])
    for r in range(X.shape[0]):
        inside = np.all((X[r] >= boxes[:, :, 0]) & (X[r] <= boxes[:, :, 1]), axis=1)
        k = np.flatnonzero(inside)
        if len(k) and r % 3 != 0:
            status[k[0]] = rng.randint(1, 4)
    extra = rng.choice(len(p_list), 200, replace=False)
    status[extra] = np.where(status[extra] == 0, rng.randint(1, 4, size=200), status[extra])
    for k in np.flatnonzero(status):
        partition_results[ns["partition_to_key"](p_list[k])] = names[int(status[k])]
    import collections
    counters = collections.defaultdict(int)
    preds, logits, routes = [], [], []
    for r in range(X.shape[0]):
        st, key = ns["find_partition_result_for_point"](X[r], partition_results)
        pred = ns["hybrid_predict"](X[r], orig, fair, partition_results, counters, "AC-3", "AC-1")
        used = fair if st == "sat" else orig
        preds.append(np.float32(pred)); logits.append(used.last_logit)
        routes.append({None: 0, "sat": 1, "unsat": 2, "unknown": 3}[st])
    out = {"X": X, "status": status, "T": np.int64(T), "pred": np.array(preds, dtype=np.float32),
           "logit": np.array(logits, dtype=np.float32), "route": np.array(routes, dtype=np.uint8),
           "counters": np.array([counters["fallback_to_original"], counters["sat_unfair_ac-1_used"],
                                 counters["unsat_fair_ac-3_used"], counters["unknown_ac-3_used"]], dtype=np.int64)}
    np.savez_compressed(os.path.join(HERE, "hybrid_case.npz"), **out)
    print("hybrid case: routes", np.bincount(out["route"], minlength=4), "counters", out["counters"])


def relaxed_case():
    """The relaxed fairness property exactly as relaxed/AC/Verify-AC.py:122-132 assembles it for one partition --
    in_const_adult ('neq' on the protected attribute, 'eq' on the others), in_const_diff_adult on the relaxed attribute,
    in_const_domain_adult on the partition (utils/verif_utils.py:1065-1091, :859-867, :781-799) -- evaluated by Z3 on
    concrete pairs of points: python tests/golden/make_golden.py relaxed  ->  tests/golden/relaxed_case.npz"""
    import importlib
    import z3
    ref = refshim.load()
    vu = importlib.import_module("utils.verif_utils")
    D = domains.DOMAINS["AC"]
    df = pd.DataFrame(np.zeros((2, D.d + 1)), columns=D.columns + ["income-per-year"])
    rd = D.copy_ranges()                                                  # the ranges of relaxed/AC/Verify-AC.py:36-48
    PA, RA, thr = ["race"], ["age"], 5                                    # :50-53
    p_list = ref.input_partition.partitioned_ranges(rd.keys(), PA, ref.input_partition.partition(rd, 6), rd)
    rng = np.random.RandomState(33)
    boxes, X, X2, ok = [], [], [], []
    x = np.array([z3.Int("x%s" % i) for i in range(D.d)]); x_ = np.array([z3.Int("x_%s" % i) for i in range(D.d)])
    pa, ra = D.columns.index("race"), D.columns.index("age")
    for pidx in rng.choice(len(p_list), 6, replace=False):
        p = p_list[int(pidx)]
        props = []
        for attr in rd.keys():
            if attr in PA:
                props.extend(vu.in_const_adult(df, x, attr, "neq", x_))
            elif attr in RA:
                props.extend(vu.in_const_diff_adult(df, x, x_, attr, thr))
            else:
                props.extend(vu.in_const_adult(df, x, attr, "eq", x_))
        props.extend(vu.in_const_domain_adult(df, x, x_, p, PA))
        prop = z3.And(*props)
        box = box_array(p, D.columns)
        for k in range(120):
            a = rng.randint(box[:, 0], box[:, 1] + 1)
            if k % 9 == 0:                                                # first point outside the partition
                c = rng.randint(D.d); a[c] = box[c, 1] + 1 + rng.randint(3)
            b = a.copy()
            b[pa] = rng.randint(-1, 7)                                    # protected value: same, other, or out of range
            b[ra] = a[ra] + rng.randint(-7, 8)                            # relaxed attribute: within or beyond the threshold
            if k % 5 == 0:                                                # some other attribute differs
                c = rng.choice([i for i in range(D.d) if i not in (pa, ra)]); b[c] += 1
            sub = [(x[i], z3.IntVal(int(a[i]))) for i in range(D.d)] + [(x_[i], z3.IntVal(int(b[i]))) for i in range(D.d)]
            val = z3.simplify(z3.substitute(prop, *sub))
            assert z3.is_true(val) or z3.is_false(val)
            boxes.append(box); X.append(a); X2.append(b); ok.append(z3.is_true(val))
    out = {"box": np.array(boxes, dtype=np.int32), "x": np.array(X, dtype=np.int64), "x2": np.array(X2, dtype=np.int64),
           "admissible": np.array(ok, dtype=np.uint8), "pa": np.int64(pa), "ra": np.int64(ra), "thr": np.int64(thr)}
    np.savez_compressed(os.path.join(HERE, "relaxed_case.npz"), **out)
    print("relaxed case:", len(ok), "pairs,", int(np.sum(ok)), "admissible")


def counterexample_cases():
    """Known answers shipped with the reference: experimentData/task5/AC/counterexamples-fairify-AC-{1,2,3}.csv, the decoded
    counterexample pairs its drivers wrote (src/AC/Verify-AC-experiment-new.py:344-399) with `output` = sigmoid(net(x)) of
    the float32-cast point and `prediction` = its class.  The rows are re-encoded with the encoders load_adult_ac1 fits
    (utils/verif_utils.py:133-171: LabelEncoder per categorical column, 20 uniform bins for capital-gain / -loss, decoded
    as the bin's integer midpoint) on data/adult/adult.{data,test}:
    python tests/golden/make_golden.py counterexamples  ->  tests/golden/counterexamples.npz"""
    from sklearn.preprocessing import KBinsDiscretizer, LabelEncoder
    cols = ['age', 'workclass', 'fnlwgt', 'education', 'education-num', 'marital-status', 'occupation', 'relationship',
            'race', 'sex', 'capital-gain', 'capital-loss', 'hours-per-week', 'native-country', 'income-per-year']
    train = pd.read_csv(os.path.join(REF, "data", "adult", "adult.data"), header=None, names=cols, skipinitialspace=True, na_values=['?'])
    test = pd.read_csv(os.path.join(REF, "data", "adult", "adult.test"), header=0, names=cols, skipinitialspace=True, na_values=['?'])
    df = pd.concat([test, train], ignore_index=True).drop(columns=['fnlwgt']).dropna()
    enc = {}
    for f in ['sex', 'workclass', 'education', 'marital-status', 'occupation', 'relationship', 'native-country', 'race']:
        enc[f] = LabelEncoder().fit(df[f])
    for f in ['capital-gain', 'capital-loss']:
        enc[f] = KBinsDiscretizer(n_bins=20, encode='ordinal', strategy='uniform').fit(df[[f]])
    D = domains.DOMAINS["AC"]
    weights = np.load(os.path.join(HERE, "weights.npz"))
    out = {}
    for name in ("AC-1", "AC-2", "AC-3"):
        c = pd.read_csv(os.path.join(REF, "experimentData", "task5", "AC", f"counterexamples-fairify-{name}.csv"))
        X = []
        for _, r in c.iterrows():
            x = []
            for f in D.columns:
                if isinstance(enc.get(f), LabelEncoder):
                    x.append(int(enc[f].transform([r[f]])[0]))
                elif f in enc:
                    e = enc[f].bin_edges_[0]
                    x.append([int((e[k] + e[k + 1]) / 2) for k in range(len(e) - 1)].index(int(r[f])))
                else:
                    x.append(int(r[f]))
            X.append(x)
        X = np.array(X, dtype=np.int64)
        # the re-encoding is right if the reference's own arithmetic (float32 weights on the float32-cast point) gives the column back
        L = int(weights[f"{name}/L"])
        a = X.astype(np.float32)
        for l in range(L):
            a = a @ weights[f"{name}/w{l}"] + weights[f"{name}/b{l}"]
            if l < L - 1:
                a = np.maximum(a, np.float32(0))
        got = 1.0 / (1.0 + np.exp(-a[:, 0].astype(np.float64)))
        assert np.allclose(got, c["output"].to_numpy(), rtol=0, atol=5e-6), name
        assert np.array_equal((got > 0.5).astype(int), c["prediction"].to_numpy()), name
        lo = np.array([D.ranges[f][0] for f in D.columns]); hi = np.array([D.ranges[f][1] for f in D.columns])
        assert np.all((X >= lo) & (X <= hi)), name
        out[f"{name}/x"] = X; out[f"{name}/output"] = c["output"].to_numpy(); out[f"{name}/prediction"] = c["prediction"].to_numpy().astype(np.uint8)
        print(name, len(X), "rows, max |sigmoid - output|", np.abs(got - c["output"].to_numpy()).max())
    np.savez_compressed(os.path.join(HERE, "counterexamples.npz"), **out)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "hybrid":
        hybrid_case()
    elif len(sys.argv) > 1 and sys.argv[1] == "singular":
        singular_cases()
    elif len(sys.argv) > 1 and sys.argv[1] == "sliver":
        sliver_case()
    elif len(sys.argv) > 1 and sys.argv[1] == "relaxed":
        relaxed_case()
    elif len(sys.argv) > 1 and sys.argv[1] == "counterexamples":
        counterexample_cases()
    else:
        main()
