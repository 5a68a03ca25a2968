"""ctypes binding of libb200knn.so (the C ABI in include/b200knn.h).

The library is built in-tree by ``make lib`` (or ``__graft_entry__.build()``) into
``build/libb200knn.so``; nothing is JIT-compiled at import time. On a machine with a GPU a missing
library is a hard error: there is no silent PyTorch fallback for any op in this package.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

REPO_ROOT = Path(__file__).resolve().parents[2]
LIB_PATH = REPO_ROOT / "build" / "libb200knn.so"

IPC_BLOB_BYTES = 256
MAX_RANKS = 16


class StepArgs(C.Structure):
    """Mirror of ``b200knn_step_args`` (include/b200knn.h)."""

    _fields_ = [
        ("k", C.c_int), ("world", C.c_int), ("rank", C.c_int),
        ("groups_per_chunk", C.c_int), ("num_chunks", C.c_int),
        ("step_begin", C.c_int), ("step_end", C.c_int),
        ("max_radius", C.c_float),
        ("inbox_heap", C.c_void_p * 2), ("inbox_pos", C.c_void_p * 2), ("inbox_cnt", C.c_void_p * 2),
        ("outbox_heap", C.c_void_p * 2), ("outbox_pos", C.c_void_p * 2), ("outbox_cnt", C.c_void_p * 2),
        ("result_sorted", C.c_void_p),
    ]


_lib = None


def build_library(force: bool = False) -> Path:
    """Compile the library with nvcc for sm_100a (cross-compiles without a GPU)."""
    if force or not LIB_PATH.exists():
        subprocess.run(["make", "-C", str(REPO_ROOT), "all"], check=True,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    return LIB_PATH


def _proto(lib):
    vp, i32, i64, f32 = C.c_void_p, C.c_int, C.c_int64, C.c_float
    P = C.POINTER
    lib.b200knn_last_error.restype = C.c_char_p
    lib.b200knn_last_error.argtypes = []
    lib.b200knn_version.restype = i32
    lib.b200knn_launch_count.restype = C.c_uint64
    lib.b200knn_launch_count.argtypes = []
    sig = {
        "b200knn_index_create": [vp, i64, vp, C.c_uint32, vp, P(vp)],
        "b200knn_index_bounds": [vp, vp],
        "b200knn_index_copy_sorted": [vp, vp],
        "b200knn_index_copy_boxes": [vp, i32, vp, P(i32)],
        "b200knn_knn_self": [vp, i32, f32, vp, vp],
        "b200knn_knn_query": [vp, vp, i64, i32, f32, vp, vp],
        "b200knn_fcp_query": [vp, vp, i64, f32, vp, vp, vp],
        "b200knn_index_stats": [vp, vp],
        "b200knn_index_enable_stats": [vp, i32],
        "b200knn_ring_create": [vp, i32, i32, i32, i64, i32, P(vp)],
        "b200knn_ring_disconnect": [vp],
        "b200knn_ring_export": [vp, vp],
        "b200knn_ring_connect": [vp, vp],
        "b200knn_ring_run": [vp, f32, vp],
        "b200knn_ring_results": [vp, vp, vp],
        "b200knn_run_steps": [vp, P(StepArgs), vp],
        "b200knn_index_unsort": [vp, vp, vp, vp],
        "b200knn_cpu_bruteforce": [vp, i64, vp, i64, i32, f32, vp, i32],
        "b200knn_cpu_knn": [vp, i64, vp, i64, i32, f32, vp, i32],
    }
    for name, args in sig.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = i32
    lib.b200knn_index_destroy.argtypes = [vp]
    lib.b200knn_index_destroy.restype = None
    lib.b200knn_ring_destroy.argtypes = [vp]
    lib.b200knn_ring_destroy.restype = None
    lib.b200knn_index_num_points.argtypes = [vp]
    lib.b200knn_index_num_points.restype = i64
    lib.b200knn_index_num_levels.argtypes = [vp]
    lib.b200knn_index_num_levels.restype = i32
    lib.b200knn_ring_comm_bytes.argtypes = [vp]
    lib.b200knn_ring_comm_bytes.restype = i64
    return lib


def lib():
    """Load (building first if the .so is absent) and return the ctypes handle."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            if os.environ.get("B200KNN_NO_AUTOBUILD"):
                raise RuntimeError(f"{LIB_PATH} is missing; run `make lib`")
            build_library()
        _lib = _proto(C.CDLL(str(LIB_PATH)))
    return _lib


class KnnError(RuntimeError):
    pass


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().b200knn_last_error().decode("utf-8", "replace")
        raise KnnError(f"{what or 'b200knn'} failed (rc={rc}): {msg}")
