"""`layer_net`, `net`, `z3_net` for any model name (reference utils/<MODEL>-Model-Functions.py, e.g.
utils/AC-1-Model-Functions.py:16-48; the 53 files differ only in the unrolled depth).

`net` and `layer_net` keep the reference's per-sample signatures; a single call is one forward on the device.
The drivers call them a handful of times per partition (C-check, src/AC/Verify-AC.py:229-249); the bulk
paths (`get_y_pred`, `candidate_dead_nodes`) recognise these functions and batch instead of looping.
"""
import numpy as np

from utils import _device


def layer_net(x, w, b):
    """Activations of every layer for one input (ReLU on all but the last layer)."""
    return _device.layer_activations(np.asarray(x), w, b)


def net(x, w, b):
    """Pre-sigmoid logit, a 1-element array like the reference's `x3` (float32 when x and w are float32)."""
    return _device.logits(np.asarray(x)[None, :], w, b)[0:1]


def z3_net(x, w, b):
    """Symbolic forward over Z3 reals (host; not on the GPU path).  Same construction as the reference:
    inputs through ToReal, W^T x + b, If(y >= 0, y, 0) between layers, last layer linear."""
    import z3
    cur = np.array([z3.ToReal(v) for v in x], dtype=object)
    for i in range(len(w)):
        pre = w[i].T @ cur + b[i]
        cur = pre if i == len(w) - 1 else np.vectorize(lambda y: z3.If(y >= 0, y, z3.RealVal(0)))(pre)
    return cur
