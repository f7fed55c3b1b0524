"""ctypes binding of liblmi_b200.so (the C-ABI declared in include/lmi_b200.h).

There is deliberately no fallback: if the shared library is missing the import raises, and if it
loads on a machine whose GPU is not sm_100 ``lmi_create`` fails with a clear message.
"""
from __future__ import annotations

import ctypes as C
import os

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG_DIR, "liblmi_b200.so")

# every symbol include/lmi_b200.h declares (tests check the library exports all of them)
SYMBOLS = [
    "lmi_last_error", "lmi_version", "lmi_create", "lmi_destroy", "lmi_set_option", "lmi_set_shard",
    "lmi_set_mlp", "lmi_store_begin", "lmi_store_append", "lmi_store_append_strided",
    "lmi_store_end", "lmi_store_save", "lmi_store_load", "lmi_set_buckets",
    "lmi_navigate", "lmi_predict", "lmi_search", "lmi_search_device", "lmi_search_rank",
    "lmi_partial_bytes", "lmi_search_partial_device", "lmi_navigate_device", "lmi_search_partial_classes_device", "lmi_merge_partials_device", "lmi_get_stats",
    "lmi_get_bucket_ptr", "lmi_rescore_kernel_name", "lmi_navigate_device_noalloc",
]


class LmiStats(C.Structure):
    _fields_ = [
        ("nq", C.c_int32), ("n_buckets", C.c_int32), ("k", C.c_int32), ("kprime", C.c_int32),
        ("n_objects_local", C.c_int64), ("n_work_items", C.c_int64), ("pairs", C.c_int64),
        ("bucket_bytes", C.c_int64), ("n_partial_lists", C.c_int64), ("n_uncertified", C.c_int64),
        ("max_abs_score_err", C.c_double),
        ("ms_navigate", C.c_float), ("ms_plan", C.c_float), ("ms_rescore", C.c_float),
        ("ms_finalize", C.c_float), ("ms_total", C.c_float),
        ("ms_rescore_avg", C.c_float), ("n_rescore_timed", C.c_int32),
        ("padded_pairs", C.c_int64), ("k3_sm_mhz", C.c_float), ("chunk_objects", C.c_int32),
    ]

    def as_dict(self):
        return {f: getattr(self, f) for f, _ in self._fields_}


_lib = None


def load() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: the CUDA extension has not been built. Run "
            "`python -c 'import __graft_entry__ as g; g.build()'` (needs nvcc). "
            "lmi_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    p, i, i64, vp = C.c_void_p, C.c_int, C.c_int64, C.c_void_p
    lib.lmi_last_error.restype = C.c_char_p
    lib.lmi_version.restype = C.c_char_p
    lib.lmi_rescore_kernel_name.restype = C.c_char_p
    lib.lmi_rescore_kernel_name.argtypes = [p]
    lib.lmi_create.argtypes = [C.POINTER(p), i, i, i, i]
    lib.lmi_destroy.argtypes = [p]
    lib.lmi_destroy.restype = None
    lib.lmi_set_option.argtypes = [p, C.c_char_p, i64]
    lib.lmi_set_shard.argtypes = [p, i, i, i]
    lib.lmi_set_mlp.argtypes = [p, i, C.POINTER(C.c_int32), C.POINTER(vp), C.POINTER(vp)]
    lib.lmi_store_begin.argtypes = [p, vp, i64, i64, i]
    lib.lmi_store_append.argtypes = [p, vp, i64, i64, i]
    lib.lmi_store_append_strided.argtypes = [p, vp, i64, i64, i64, i64, i]
    lib.lmi_store_save.argtypes = [p, C.c_char_p, C.c_char_p]
    lib.lmi_store_load.argtypes = [p, C.c_char_p, C.c_char_p]
    lib.lmi_store_end.argtypes = [p]
    lib.lmi_set_buckets.argtypes = [p, vp, vp, i64, i64]
    lib.lmi_navigate.argtypes = [p, vp, i64, i, vp, vp]
    lib.lmi_predict.argtypes = [p, vp, i64, vp]
    lib.lmi_search.argtypes = [p, vp, vp, i64, i, i, vp, vp]
    lib.lmi_search_device.argtypes = [p, vp, vp, i64, i, i, vp, vp, vp]
    lib.lmi_search_rank.argtypes = [p, vp, i64, vp, i, vp, vp, vp]
    lib.lmi_partial_bytes.argtypes = [i64, i]
    lib.lmi_partial_bytes.restype = i64
    lib.lmi_search_partial_device.argtypes = [p, vp, vp, i64, i, i, vp, vp]
    lib.lmi_navigate_device.argtypes = [p, vp, i64, i, vp, vp]
    lib.lmi_search_partial_classes_device.argtypes = [p, vp, vp, i64, i, i, vp, vp]
    lib.lmi_merge_partials_device.argtypes = [p, vp, i, i64, i, vp, vp, vp]
    lib.lmi_navigate_device_noalloc.argtypes = [p, vp, i64, i, vp, vp, vp]
    lib.lmi_get_stats.argtypes = [p, C.POINTER(LmiStats)]
    lib.lmi_get_bucket_ptr.argtypes = [p, vp]
    _lib = lib
    return lib


class LmiError(RuntimeError):
    pass


def check(rc: int):
    """Mirror the reference's error behaviour: state errors are AssertionError (LearnedIndex.py:60),
    everything else RuntimeError / ValueError."""
    if rc == 0:
        return
    msg = load().lmi_last_error().decode()
    if rc == -2:
        raise AssertionError(msg)
    if rc == -1:
        raise ValueError(msg)
    raise LmiError(f"lmi_b200 error {rc}: {msg}")
