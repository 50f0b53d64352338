"""Model cache and thin helpers between the reference-shaped API and fairify_b200.engine."""
import collections
import hashlib

import numpy as np

from fairify_b200 import engine

_models = {}


_by_id = collections.OrderedDict()


def model_for(w, b):
    """engine.Model for these weights (cached by content: the drivers pass freshly pruned arrays each time; and by
    identity of the arrays, which are kept alive while cached, so that the partition loop's repeated calls with the
    same lists do not re-hash the weights)."""
    ident = tuple(id(a) for a in w) + tuple(id(a) for a in b)
    hit = _by_id.get(ident)
    if hit is not None:
        return hit[0]
    m = _model_by_content(w, b)
    _by_id[ident] = (m, list(w), list(b))
    while len(_by_id) > 16:
        _by_id.popitem(last=False)
    return m


def _model_by_content(w, b):
    h = hashlib.blake2b(digest_size=16)
    for a in list(w) + list(b):
        a = np.ascontiguousarray(a, dtype=np.float32)
        h.update(str(a.shape).encode()); h.update(a.tobytes())
    key = h.digest()
    m = _models.get(key)
    if m is None:
        if len(_models) > 64:
            _models.clear()
        m = engine.Model([np.asarray(x, dtype=np.float32) for x in w], [np.asarray(x, dtype=np.float32) for x in b])
        _models[key] = m
    return m


# Pruned networks.  The drivers call prune_neurons() once per partition and then predict with the compacted weights
# (get_y_pred on the 1000 samples, the C-check: src/AC/Verify-AC.py:229-274).  Compaction only deletes terms that are
# zero anyway, so the pruned network equals the original one with the dead neurons' activations forced to 0 -- which
# fb_forward_* does with a mask on the resident model.  prune_neurons registers its result here (the arrays are kept
# alive while registered, so their ids cannot be reused); logits() recognises them.
_pruned = collections.OrderedDict()
_PRUNED_KEEP = 8


def remember_pruned(pr_w, pr_b, weight, bias, dead_lists):
    key = tuple(id(a) for a in pr_w) + tuple(id(a) for a in pr_b)
    mask = np.concatenate([np.asarray(l, dtype=np.uint8) != 0 for l in dead_lists]).astype(np.uint8)
    _pruned[key] = (list(pr_w), list(pr_b), list(weight), list(bias), mask)
    while len(_pruned) > _PRUNED_KEEP:
        _pruned.popitem(last=False)


def _as_pruned(w, b):
    """(original model, dead mask words) when (w, b) is a registered prune_neurons result, else None."""
    hit = _pruned.get(tuple(id(a) for a in w) + tuple(id(a) for a in b))
    if hit is None:
        return None
    _, _, weight, bias, mask = hit
    m = model_for(weight, bias)
    return m, engine.pack_bits(mask[None])[0]


def logits(X, w, b):
    """numpy logits for rows X: float32 when both X and the weights are float32 (numpy's promotion in
    `w.T @ x`), float64 otherwise."""
    X = np.asarray(X)
    f32 = X.dtype == np.float32 and all(np.asarray(a).dtype == np.float32 for a in w)
    X = X.astype(np.float32 if f32 else np.float64)
    pruned = _as_pruned(w, b)
    if pruned is not None:
        m, dead_bits = pruned
        return m.forward(X, dead_bits=dead_bits).cpu().numpy()
    return model_for(w, b).forward(X).cpu().numpy()


def layer_activations(x, w, b):
    """Per-layer activations of one input.  Only the interactive `layer_net` call needs them; they are the
    device bounds of the degenerate box [x, x] (lower == upper == the pre-activation)."""
    x = np.asarray(x)
    if np.issubdtype(x.dtype, np.integer):
        m = model_for(w, b)
        box = np.stack([x, x], axis=1).astype(np.int32)[None]
        # degenerate box: lower == upper == W^T x + b summed in the reference order of neuron_bounds
        r = m.ibp(box, want=("ws_lb",))
        pre = r.numpy("ws_lb")[0]
        acts = m.split(np.where(pre > 0, pre, 0.0))
        acts[-1] = m.split(pre)[-1]
        return [np.array(a) for a in acts]
    raise TypeError("layer_net is only defined for integer sample rows on the GPU path "
                    "(the reference calls it from candidate_dead_nodes alone, utils/prune.py:169,176)")
