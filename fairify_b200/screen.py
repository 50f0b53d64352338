"""Sampled protected-attribute pair screen on the GPU.

Mirror of `CausalDiscriminationDetector` (reference `src/AC/metrics.py:40-264`, identical copy in
`src/GC/metrics.py`).  The reference draws one random input at a time, flips the protected attributes
through every value, predicts each variant with one keras `model.predict` call, and stops on a confidence
interval.  Its random stream (`random.seed(42)`, `random.choice`) never depends on a prediction
(`:132-135`), so the draws are replayed up front with the same `random.Random(42)` stream, all
`max_samples-1` base rows are screened in ONE kernel launch (`fb_pair_screen`), and the counting /
stopping rule (`:161-166`, `:243-257`) is replayed on the host.  Results are identical to the reference's
for the same float32 predictions.
"""
from __future__ import annotations

import itertools
import math
import random

import numpy as np


def _z_score(conf):
    import scipy.stats as st          # what the reference calls (:254)
    return st.norm.ppf(conf)


def draw_base_rows(columns, values, protected, max_samples=1000, random_seed=42):
    """The rows `causal_discrimination` would draw: float32 [max_samples-1, d].

    `values[c]` is the value list of column c in the order given to `add_feature`; fixed features are drawn
    first in input order, then the protected ones in the order of `protected` (:132-135).
    """
    rng = random.Random(random_seed)
    fixed = [c for c in columns if c not in protected]
    col = {c: i for i, c in enumerate(columns)}
    X = np.zeros((max_samples - 1, len(columns)), dtype=np.float32)
    for r in range(max_samples - 1):
        for c in fixed:
            X[r, col[c]] = rng.choice(values[c])
        for c in protected:
            X[r, col[c]] = rng.choice(values[c])
    return X


def causal_discrimination(model, columns, values, protected, conf=0.999, margin=0.0001, max_samples=1000,
                          min_samples=100, random_seed=42):
    """-> (test_cases, discrimination_rate, causal_pairs), as `causal_discrimination` (:101-168) returns
    them, with every case a list of floats in column order instead of a dict of strings."""
    assert protected, "Must specify protected features to test"
    columns = list(columns)
    X = draw_base_rows(columns, values, protected, max_samples, random_seed)
    pa_cols = [columns.index(c) for c in protected]
    alts = np.array(list(itertools.product(*[values[c] for c in protected])), dtype=np.float32).reshape(-1, len(protected))
    orig, flip = model.pair_screen(X, pa_cols, alts)
    flip = flip.cpu().numpy()
    count = 0
    test_cases, pairs = [], []
    rate = 0
    for r in range(max_samples - 1):
        num_sampled = r + 1
        base = X[r]
        test_cases.append(base.tolist())
        own = base[pa_cols]
        for a in range(alts.shape[0]):
            if np.array_equal(alts[a], own):
                continue
            mod = base.copy(); mod[pa_cols] = alts[a]
            test_cases.append(mod.tolist())
            if a == flip[r]:
                count += 1
                pairs.append((base.tolist(), mod.tolist()))
                break
        if num_sampled < min_samples:
            rate, stop = 0, False
        else:
            rate = count / num_sampled
            if rate == 0 or rate == 1:
                error = 0
            else:
                error = _z_score(conf) * math.sqrt((rate * (1 - rate)) / num_sampled)
            stop = error < margin
        if stop:
            break
    return test_cases, rate, pairs
