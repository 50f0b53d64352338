"""Optional bridge to a reference checkout for everything that is NOT on the GPU path.

The Z3 property builders (`in_const_*`, `in_const_domain_*`, `in_const_diff_*`, `parse_z3Model`), the dataset
loaders and friends stay host code exactly as in the reference (north star: "the drivers and Z3 encoding
stay").  When FAIRIFY_REFERENCE_ROOT (default /root/reference) holds a checkout, its utils/verif_utils.py is
loaded under a private name and the names this package does not define are re-exported from it, so the
unchanged drivers find everything they star-import.  tensorflow / aif360 are stubbed if absent (only
`load_model` and `StandardDataset` are touched at import time).
"""
import contextlib
import importlib.util
import io
import os
import sys
import types

ROOT = os.environ.get("FAIRIFY_REFERENCE_ROOT", "/root/reference")
_cache = {}


def available():
    return os.path.isfile(os.path.join(ROOT, "utils", "verif_utils.py"))


def _stub(name, **attrs):
    if name in sys.modules:
        return sys.modules[name]
    m = types.ModuleType(name)
    for k, v in attrs.items():
        setattr(m, k, v)
    sys.modules[name] = m
    return m


def load(name):
    """Reference utils/<name>.py as a private module, or None."""
    if name in _cache:
        return _cache[name]
    mod = None
    if available():
        try:
            def _no_tf(*a, **k):
                raise RuntimeError("tensorflow is not installed")
            try:
                import tensorflow  # noqa: F401
            except Exception:
                tf = _stub("tensorflow"); keras = _stub("tensorflow.keras")
                tf.keras = keras
                keras.models = _stub("tensorflow.keras.models", load_model=_no_tf, Model=object)
                keras.backend = _stub("tensorflow.keras.backend")
            try:
                import aif360  # noqa: F401
            except Exception:
                _stub("aif360").datasets = _stub("aif360.datasets", StandardDataset=object)
            if name != "standard_data" and "utils.standard_data" not in sys.modules:
                sd = load("standard_data")
                if sd is not None:
                    sys.modules["utils.standard_data"] = sd
            spec = importlib.util.spec_from_file_location(f"_fairify_reference_{name}", os.path.join(ROOT, "utils", f"{name}.py"))
            mod = importlib.util.module_from_spec(spec)
            with contextlib.redirect_stdout(io.StringIO()):
                spec.loader.exec_module(mod)
        except Exception:
            mod = None
    _cache[name] = mod
    return mod


def reexport(name, namespace, skip=()):
    """Copy the reference module's public names that `namespace` does not define yet."""
    mod = load(name)
    if mod is None:
        return []
    added = []
    for k, v in vars(mod).items():
        if k.startswith("_") or k in namespace or k in skip:
            continue
        namespace[k] = v
        added.append(k)
    return added
