"""ctypes binding of libfairify_b200.so (C ABI: include/fairify_b200.h).

Loading fails loudly when the library has not been built; compute calls fail loudly (RuntimeError with
fb_last_error()) when no CUDA device is usable.  There is no CPU path behind this module.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.environ.get("FAIRIFY_B200_LIB") or os.path.join(_HERE, "libfairify_b200.so")

# every symbol include/fairify_b200.h declares
EXPORTS = [
    "fb_version", "fb_last_error", "fb_num_sms", "fb_model_create", "fb_model_destroy", "fb_model_info", "fb_model_device",
    "fb_release_cache",
    "fb_grid_create", "fb_grid_destroy", "fb_grid_size", "fb_grid_boxes", "fb_grid_locate", "fb_profile", "fb_ibp", "fb_prune",
    "fb_prune_host", "fb_prune_noalloc", "fb_forward_f64", "fb_forward_f32", "fb_pair_screen", "fb_partition_screen", "fb_heuristic", "fb_fp64_peak",
    "fb_partition_screen_ex", "fb_hybrid_predict", "fb_profile_nearzero",
]


class PruneOut(C.Structure):
    """struct fb_prune_out"""
    _fields_ = [(k, C.c_void_p) for k in ("cand", "b_dead", "dead", "sliver", "verdict", "counts",
                                          "ws_lb", "ws_ub", "pl_lb", "pl_ub", "lb_rd", "ub_ru",
                                          "s_dead", "s_cand", "record")]


class FairifyB200Error(RuntimeError):
    pass


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise FairifyB200Error(
            f"{SO_PATH} is missing: build it with `python -m fairify_b200.build` (nvcc, sm_100a). "
            "The pre-solver stage has no CPU fallback.")
    L = C.CDLL(SO_PATH)
    vp, i32, i64, u64 = C.c_void_p, C.c_int, C.c_int64, C.c_uint64
    L.fb_version.restype = i32
    L.fb_last_error.restype = C.c_char_p
    L.fb_num_sms.argtypes = [C.POINTER(i32)]
    L.fb_model_create.argtypes = [i32, vp, vp, vp, C.POINTER(vp)]
    L.fb_model_destroy.argtypes = [vp]; L.fb_model_destroy.restype = None
    L.fb_model_info.argtypes = [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)]
    L.fb_model_device.argtypes = [vp, C.POINTER(i32)]
    L.fb_release_cache.argtypes = []
    L.fb_grid_create.argtypes = [i32, vp, vp, i64, vp, C.POINTER(vp)]
    L.fb_grid_destroy.argtypes = [vp]; L.fb_grid_destroy.restype = None
    L.fb_grid_size.argtypes = [vp]; L.fb_grid_size.restype = i64
    L.fb_grid_boxes.argtypes = [vp, i64, i64, vp, vp]
    L.fb_grid_locate.argtypes = [vp, i64, vp, vp, vp, vp]
    L.fb_profile.argtypes = [vp, i64, vp, vp, i32, u64, i64, vp, vp]
    L.fb_ibp.argtypes = [vp, i64, vp, vp, vp, C.POINTER(PruneOut), vp]
    L.fb_prune.argtypes = [vp, i64, vp, vp, i32, u64, i64, vp, C.POINTER(PruneOut), vp]
    L.fb_prune_host.argtypes = [vp, i64, vp, vp, i32, u64, i64, vp, C.POINTER(PruneOut), vp]
    L.fb_prune_noalloc.argtypes = [vp, i64, vp, vp, i32, u64, i64, vp, C.POINTER(PruneOut), vp]
    L.fb_forward_f64.argtypes = [vp, i64, vp, vp, vp, vp]
    L.fb_forward_f32.argtypes = [vp, i64, vp, vp, vp, vp]
    L.fb_pair_screen.argtypes = [vp, i64, vp, i32, vp, i32, vp, vp, vp, vp]
    L.fb_partition_screen.argtypes = [vp, i64, vp, vp, i32, u64, i64, i32, vp, vp, vp, vp, vp]
    L.fb_partition_screen_ex.argtypes = [vp, i64, vp, vp, i32, u64, i64, i32, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp, vp]
    L.fb_hybrid_predict.argtypes = [vp, vp, vp, i64, vp, vp, vp, vp, vp, vp, vp]
    L.fb_profile_nearzero.argtypes = [vp, i64, vp, vp, i32, u64, i64, vp, vp, vp]
    L.fb_heuristic.argtypes = [vp, i64, vp, vp, vp, vp, vp, C.c_double, vp, vp, vp]
    L.fb_fp64_peak.argtypes = [i32, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    _lib = L
    return L


def check(rc, what=""):
    if rc != 0:
        msg = lib().fb_last_error().decode("utf-8", "replace")
        raise FairifyB200Error(f"{what}: {msg}" if what else msg)
