"""ctypes binding of libflagg.so (include/flagg.h).  No fallback: if the shared library
is missing or a call fails, an exception is raised -- the reference's Flower server
swallows defense exceptions into plain FedAvg (robust_server.py:120-122), so a silent
fallback here would mask a broken kernel."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FLAGG_LIB") or os.path.join(_HERE, "libflagg.so")   # FLAGG_LIB: tuning builds only

MEDIAN, TRIMMED_MEAN = 0, 1
STATIC_SCRATCH_BYTES = 8 << 20          # FLAGG_STATIC_SCRATCH_BYTES (flagg_row_reduce_static)
T_SCALE, T_REPLACE, T_SIGN, T_AXPY = 0, 1, 2, 3


class FlaggError(RuntimeError):
    pass


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FlaggError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C multimodal-fl-security_b200/csrc`). There is no CPU fallback.")
    L = C.CDLL(LIB_PATH)
    vp, i32, i64, f32 = C.c_void_p, C.c_int, C.c_int64, C.c_float
    sig = {
        "flagg_version": (i32, []),
        "flagg_last_error": (C.c_char_p, []),
        "flagg_device_info": (i32, [C.POINTER(i32)] * 3),
        "flagg_view_create": (i32, [C.POINTER(vp), i32, i32, C.POINTER(i64), C.POINTER(vp), vp]),
        "flagg_view_destroy": (i32, [vp]),
        "flagg_view_set_outputs": (i32, [vp, C.POINTER(vp), vp]),
        "flagg_view_cols": (i64, [vp]),
        "flagg_view_rows": (i32, [vp]),
        "flagg_colwise_select": (i32, [vp, i32, i32, vp, vp]),
        "flagg_row_reduce": (i32, [vp, vp, i32, vp, vp, vp, vp, vp]),
        "flagg_row_reduce_static": (i32, [vp, vp, i32, vp, vp, vp, vp, vp]),
        "flagg_row_combine": (i32, [vp, i32, C.POINTER(C.c_int32), C.POINTER(f32), C.POINTER(f32), f32, vp, vp]),
        "flagg_row_combine_dev": (i32, [vp, i32, vp, vp, vp, vp, vp, vp, vp, vp]),
        "flagg_weiszfeld_weights": (i32, [vp, i32, vp, vp, vp]),
        "flagg_round_weights": (i32, [i32, i32, vp, vp, vp, vp, vp, C.c_double, C.c_double, vp, vp, vp, vp, vp, vp, vp, vp]),
        "flagg_gram": (i32, [vp, vp, i32, vp, vp]),
        "flagg_krum_scores": (i32, [vp, i32, i32, vp, vp, vp]),
        "flagg_krum_scores_from_distances": (i32, [vp, i32, i32, vp, vp]),
        "flagg_sqdist": (i32, [vp, i32, vp, vp]),
        "flagg_krum_scores_from_sqdist": (i32, [vp, i32, i32, vp, vp, vp]),
        "flagg_segments_transform": (i32, [i32, C.POINTER(i64), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), i32, f32, vp]),
        "flagg_colwise_select_host": (i32, [i32, i32, C.POINTER(i64), C.POINTER(vp), i32, i32, vp, i64]),
        "flagg_upload_rows": (i32, [i32, i32, C.POINTER(i64), C.POINTER(vp), vp, i64, vp]),
        "flagg_host_release": (i32, []),
        "flagg_comm_unique_id": (i32, [vp]),
        "flagg_comm_create": (i32, [C.POINTER(vp), i32, i32, vp]),
        "flagg_comm_destroy": (i32, [vp]),
        "flagg_comm_size": (i32, [vp]),
        "flagg_comm_rank": (i32, [vp]),
        "flagg_allreduce_f64": (i32, [vp, vp, C.c_size_t, i32, vp]),
        "flagg_allreduce_f32": (i32, [vp, vp, C.c_size_t, i32, vp]),
        "flagg_comm_scatter_columns": (i32, [vp, i32, i32, i32, C.POINTER(i64), C.POINTER(vp), C.POINTER(i64), C.POINTER(i64), vp, i64, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)          # AttributeError here = header/library mismatch
        fn.restype, fn.argtypes = res, args
    _lib = L
    return L


EXPORTS = ["flagg_version", "flagg_last_error", "flagg_device_info", "flagg_view_create", "flagg_view_destroy",
           "flagg_view_set_outputs", "flagg_row_combine_dev", "flagg_weiszfeld_weights", "flagg_krum_scores_from_distances", "flagg_segments_transform",
           "flagg_sqdist", "flagg_krum_scores_from_sqdist", "flagg_round_weights", "flagg_view_cols", "flagg_view_rows", "flagg_colwise_select", "flagg_row_reduce", "flagg_row_combine",
           "flagg_gram", "flagg_krum_scores", "flagg_colwise_select_host", "flagg_upload_rows", "flagg_host_release", "flagg_comm_unique_id",
           "flagg_comm_create", "flagg_comm_destroy", "flagg_comm_size", "flagg_comm_rank", "flagg_allreduce_f64",
           "flagg_allreduce_f32", "flagg_comm_scatter_columns", "flagg_row_reduce_static"]


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = lib().flagg_last_error().decode("utf-8", "replace")
        raise FlaggError(f"{what} failed (status {rc}): {msg}")
