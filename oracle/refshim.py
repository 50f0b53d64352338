"""Import the reference's hot-path modules unmodified, in the build container only.

TEST INFRASTRUCTURE ONLY.  `/root/reference` does not exist on the GPU box, so nothing that runs there
may import this module; it is used by tests/golden/make_golden.py (fixture generation) and by the
optional live cross-check tests, which skip when the reference checkout is absent.

The reference needs tensorflow and aif360 (`requirements.txt:1-3`), neither of which is installed.
Only `load_model` / `StandardDataset` are touched at import time, so empty stub modules are enough to
import `utils/prune.py`, `utils/input_partition.py` and the model-function modules.
The process that calls `load()` must not have imported this repo's own top-level `utils` package.
"""
from __future__ import annotations

import contextlib
import importlib
import io
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("FAIRIFY_REFERENCE_ROOT", "/root/reference")


def available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "utils"))


def _stub(name, **attrs):
    m = types.ModuleType(name)
    for k, v in attrs.items():
        setattr(m, k, v)
    sys.modules[name] = m
    return m


def load():
    """Return a namespace with the reference's prune / input_partition modules."""
    if not available():
        raise RuntimeError(f"reference checkout not found at {REFERENCE_ROOT}")
    if "utils" in sys.modules and not getattr(sys.modules["utils"], "__file__", "").startswith(REFERENCE_ROOT):
        raise RuntimeError("this process already imported another `utils` package")

    def _no_tf(*a, **k):
        raise RuntimeError("tensorflow is not available in this image")

    if "tensorflow" not in sys.modules:
        tf = _stub("tensorflow")
        keras = _stub("tensorflow.keras")
        models = _stub("tensorflow.keras.models", load_model=_no_tf, Model=object)
        backend = _stub("tensorflow.keras.backend")
        tf.keras = keras; keras.models = models; keras.backend = backend
    if "aif360" not in sys.modules:
        aif = _stub("aif360")
        ds = _stub("aif360.datasets", StandardDataset=object)
        aif.datasets = ds
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    if "utils" not in sys.modules:
        # the reference's utils/ has no __init__.py (a namespace package), so this repo's regular `utils` package
        # would win the path scan wherever it sits on sys.path: bind the name to the reference directory explicitly
        pkg = types.ModuleType("utils")
        pkg.__path__ = [os.path.join(REFERENCE_ROOT, "utils")]
        pkg.__file__ = os.path.join(REFERENCE_ROOT, "utils", "__init__.py")
        sys.modules["utils"] = pkg
    with contextlib.redirect_stdout(io.StringIO()):
        prune = importlib.import_module("utils.prune")
        part = importlib.import_module("utils.input_partition")
        verif = importlib.import_module("utils.verif_utils")
    return types.SimpleNamespace(prune=prune, input_partition=part, verif_utils=verif)


def model_functions(model_name):
    """`layer_net`, `net` of utils/<model>-Model-Functions.py."""
    with contextlib.redirect_stdout(io.StringIO()):
        mod = importlib.import_module(f"utils.{model_name}-Model-Functions")
    return mod.layer_net, mod.net


def metrics_module(dataset="AC"):
    """src/<dataset>/metrics.py (CausalDiscriminationDetector); needs only random + scipy."""
    path = os.path.join(REFERENCE_ROOT, "src", dataset, "metrics.py")
    spec = importlib.util.spec_from_file_location(f"ref_metrics_{dataset}", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@contextlib.contextmanager
def quiet():
    with contextlib.redirect_stdout(io.StringIO()):
        yield
