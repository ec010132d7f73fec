"""ctypes binding of libnexus_flat.so (the C ABI declared in include/nexus_flat.h).

There is deliberately no fallback: if the shared library is missing or fails to load, importing the
compute entry points raises.  Build it with ``python -m nexus_b200.build``.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

HERE = os.path.dirname(os.path.abspath(__file__))
# NEXUS_FLAT_LIB selects another build of the same C ABI (A/B timing of kernel variants on one box); default = in-tree build
LIB_PATH = os.environ.get("NEXUS_FLAT_LIB") or os.path.join(HERE, "libnexus_flat.so")

# enums (mirror include/nexus_flat.h)
NXF_OK, NXF_EINVAL, NXF_ECUDA, NXF_ENOMEM, NXF_ENODEVICE, NXF_EUNSUPPORTED = 0, -1, -2, -3, -4, -5
METRIC_IP, METRIC_L2 = 0, 1
F32, F16, BF16 = 0, 1, 2
POOL_CLS, POOL_MEAN, POOL_LAST, POOL_LAST_ARGMAX = 0, 1, 2, 3
SEARCH_DEFAULT = 0
SEARCH_FORCE_EXACT = 1 << 0
SEARCH_FORCE_STREAM = 1 << 1
SEARCH_FORCE_TENSOR = 1 << 2
SEARCH_NO_RESCORE = 1 << 3
SEARCH_TENSOR_1CTA = 1 << 4
SEARCH_FORCE_TSTREAM = 1 << 5
STREAM_MAX_NQ = 4
MAX_K = 2048
PATH_NONE, PATH_STREAM, PATH_TENSOR, PATH_EXACT, PATH_TSTREAM, PATH_SUBSET = 0, 1, 2, 3, 4, 5
PATH_NAMES = {0: "none", 1: "stream", 2: "tensor", 3: "exact", 4: "tstream", 5: "subset"}

EXPORTS = [
    "nxf_version", "nxf_last_error", "nxf_device_count", "nxf_sample_rank", "nxf_create", "nxf_destroy", "nxf_reset", "nxf_reserve",
    "nxf_ntotal", "nxf_dim", "nxf_metric", "nxf_store_dtype", "nxf_device", "nxf_set_id_offset", "nxf_get_stats",
    "nxf_set_profiling",
    "nxf_add", "nxf_reconstruct", "nxf_search", "nxf_search_subsets", "nxf_merge_topk", "nxf_positive_rank", "nxf_pool_normalize",
    "nxf_search_packed", "nxf_merge_packed", "nxf_pool_mean_normalize_static",
]


class Stats(C.Structure):
    _fields_ = [
        ("kernel_launches", C.c_int64), ("searches", C.c_int64), ("rows_added", C.c_int64),
        ("last_path", C.c_int32), ("last_k_fetch", C.c_int32), ("last_chunks", C.c_int32), ("last_grid", C.c_int32),
        ("scratch_bytes", C.c_int64), ("storage_bytes", C.c_int64),
        ("main_kernel_count", C.c_int64), ("main_kernel_ms", C.c_double),
        ("cert_checked", C.c_int64), ("cert_retry_l2", C.c_int64), ("cert_retry_l3", C.c_int64),
    ]


class NexusFlatError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"[nexus_flat {code}] {message}")
        self.code = code


_lib = None
_lock = threading.Lock()


def load() -> C.CDLL:
    """Load (once) and type the shared library.  Raises if it is not built -- no CPU fallback."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: the CUDA extension is the only compute path of nexus_b200 "
                "(no CPU fallback). Build it with `python -m nexus_b200.build`.")
        lib = C.CDLL(LIB_PATH)
        vp, i32, i64 = C.c_void_p, C.c_int, C.c_int64
        lib.nxf_version.restype = C.c_char_p
        lib.nxf_last_error.restype = C.c_char_p
        lib.nxf_device_count.restype = i32
        lib.nxf_sample_rank.argtypes = [i32, i64, i64]
        lib.nxf_create.argtypes = [i32, i32, i32, i32, C.POINTER(vp)]
        lib.nxf_destroy.argtypes = [vp]
        lib.nxf_destroy.restype = None
        lib.nxf_reset.argtypes = [vp]
        lib.nxf_reserve.argtypes = [vp, i64]
        lib.nxf_ntotal.argtypes = [vp]
        lib.nxf_ntotal.restype = i64
        for name in ("nxf_dim", "nxf_metric", "nxf_store_dtype", "nxf_device"):
            getattr(lib, name).argtypes = [vp]
            getattr(lib, name).restype = i32
        lib.nxf_set_id_offset.argtypes = [vp, i64]
        lib.nxf_get_stats.argtypes = [vp, C.POINTER(Stats)]
        lib.nxf_set_profiling.argtypes = [vp, i32]
        lib.nxf_add.argtypes = [vp, vp, i64, i32, i64, i32, vp]
        lib.nxf_reconstruct.argtypes = [vp, i64, i64, vp, i32, vp]
        lib.nxf_search.argtypes = [vp, vp, i64, i32, i64, i32, i32, vp, vp, i32, i32, vp]
        lib.nxf_search_subsets.argtypes = [vp, vp, i64, i32, i64, i32, vp, vp, i64, i32, i32, vp, vp, i32, vp]
        lib.nxf_merge_topk.argtypes = [vp, vp, i32, i64, i32, i32, i32, vp, vp, i32, vp]
        lib.nxf_search_packed.argtypes = [vp, vp, i64, i32, i64, i32, vp, i32, vp]
        lib.nxf_merge_packed.argtypes = [vp, C.POINTER(i64), i32, i64, i32, i32, i32, vp, vp, i32, vp]
        lib.nxf_positive_rank.argtypes = [vp, vp, i64, i32, vp, vp, i32, vp]
        lib.nxf_pool_normalize.argtypes = [vp, i32, vp, i64, i64, i64, i32, i32, C.c_float, vp, i32, i32, vp]
        lib.nxf_pool_mean_normalize_static.argtypes = [vp, i32, vp, i64, i64, i64, i32, C.c_float, vp, i32, i32, vp]
        for name in EXPORTS:
            fn = getattr(lib, name)  # raises AttributeError if the symbol is not exported
            if name not in ("nxf_version", "nxf_last_error", "nxf_destroy", "nxf_ntotal"):
                fn.restype = i32
        _lib = lib
    return _lib


def check(rc: int) -> int:
    if rc < 0:
        msg = load().nxf_last_error().decode("utf-8", "replace")
        if rc == NXF_EINVAL:
            raise ValueError(f"[nexus_flat {rc}] {msg}")
        raise NexusFlatError(rc, msg)
    return rc


def version() -> str:
    return load().nxf_version().decode()
