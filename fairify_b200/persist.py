"""Persist the pre-solver results of one (model, grid, sampling) run, so a rerun can skip the stage.

The reference has no checkpointing: a rerun starts over with a new shuffle (its CSV only keeps per-partition
verdicts, src/AC/Verify-AC.py:277-315).  A result file is a compressed .npz holding the packed mask / verdict /
count arrays exactly as the device wrote them, plus a fingerprint of everything they depend on (weights, grid
definition, sample count, seed, index range); `load` refuses a file whose fingerprint does not match.
"""
from __future__ import annotations

import hashlib
import json

import numpy as np

FORMAT = 2        # 2: device sampler draws 4 / 2 / 1 features per Philox word by box class (files of format 1 are refused)


def fingerprint(model, grid, S, seed, first, count):
    h = hashlib.sha256()
    for a in list(model.w) + list(model.b):
        a = np.ascontiguousarray(a, dtype=np.float32)
        h.update(str(a.shape).encode()); h.update(a.tobytes())
    meta = {"columns": list(grid.columns), "keys": list(grid.keys), "lo": grid.lo.tolist(), "hi": grid.hi.tolist(),
            "T": int(grid.T), "S": int(S), "seed": int(seed), "first": int(first), "count": int(count), "format": FORMAT}
    h.update(json.dumps(meta, sort_keys=True).encode())
    return h.hexdigest(), meta


def save(path, model, grid, S, seed, first, result):
    """`result`: engine.PruneResult (or a dict of host arrays) for canonical indices first .. first+n-1."""
    arrays = {k: (v.cpu().numpy() if hasattr(v, "cpu") else np.asarray(v)) for k, v in
              (result.t.items() if hasattr(result, "t") else result.items())}
    n = next(iter(arrays.values())).shape[0]
    fp, meta = fingerprint(model, grid, S, seed, first, n)
    np.savez_compressed(path, __fingerprint__=np.array(fp), __meta__=np.array(json.dumps(meta)), **arrays)
    return fp


def load(path, model, grid, S, seed, first, count):
    """dict of host arrays, or raises ValueError when the file belongs to another model / grid / sampling."""
    z = np.load(path, allow_pickle=False)
    fp, _ = fingerprint(model, grid, S, seed, first, count)
    if str(z["__fingerprint__"]) != fp:
        raise ValueError(f"{path}: stored results belong to a different model, grid or sampling configuration")
    return {k: z[k] for k in z.files if not k.startswith("__")}
