"""CPU: the C-ABI library loads, exports every symbol include/fairify_b200.h declares, and refuses to
compute without a CUDA device (no CPU fallback behind the product path)."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def _declared():
    text = open(os.path.join(ROOT, "include", "fairify_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(fb_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree():
    from fairify_b200 import _lib
    assert _declared() == sorted(_lib.EXPORTS)


def test_library_exports_every_declared_symbol():
    from fairify_b200 import _lib
    if not os.path.exists(_lib.SO_PATH):
        from fairify_b200 import build
        build.build()
    L = ctypes.CDLL(_lib.SO_PATH)
    for name in _declared():
        assert hasattr(L, name), name
    assert _lib.lib().fb_version() >= 100


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from fairify_b200 import engine
    from fairify_b200._lib import FairifyB200Error
    with pytest.raises(FairifyB200Error):
        engine.Model([np.zeros((3, 1), np.float32)], [np.zeros(1, np.float32)])
    # the raw ABI reports the failure through its status code, never by computing on the host
    from fairify_b200 import _lib
    h = ctypes.c_void_p()
    dims = np.array([3, 1], dtype=np.int32); w = np.zeros(3, np.float32); b = np.zeros(1, np.float32)
    rc = _lib.lib().fb_model_create(1, dims.ctypes.data_as(ctypes.c_void_p), w.ctypes.data_as(ctypes.c_void_p),
                                    b.ctypes.data_as(ctypes.c_void_p), ctypes.byref(h))
    assert rc != 0 and b"CUDA" in _lib.lib().fb_last_error()


def test_product_code_never_imports_the_oracle():
    """oracle/ is test infrastructure: nothing under fairify_b200/ or utils/ may import or load it."""
    bad = []
    for base in ("fairify_b200", "utils"):
        for dp, _, fs in os.walk(os.path.join(ROOT, base)):
            for f in fs:
                if f.endswith((".py", ".cu", ".cuh", ".h")):
                    src = open(os.path.join(dp, f), errors="replace").read()
                    if re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M) or "libfairify_oracle" in src:
                        bad.append(os.path.join(dp, f))
    assert not bad, bad
