"""CPU oracle and mirror engine (numpy in, numpy out). No GPU needed.

``bruteforce`` is the test oracle (exact fp32 expression of SURVEY.md fact 7, all-pairs,
``std::nth_element``); ``cpu_knn`` is the host restatement of the GPU algorithm. Role of
``verifyKNN`` (cudaKDTree/testing/floatN-knn-and-fcp.cu:292-335) and ``buildTree_host``
(cukd/builder_thrust.h:514-601).
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Optional

import numpy as np

from ..ops._lib import check, lib


def _f32c(a) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.float32)
    if a.ndim != 2 or a.shape[1] not in (2, 3):
        raise ValueError("expected an [N,3] (or [N,2]) array")
    if a.shape[1] == 2:                      # 2-D clouds: z = 0 gives the 2-D distance bit for bit
        a = np.ascontiguousarray(np.concatenate([a, np.zeros((len(a), 1), np.float32)], axis=1))
    return a


def bruteforce(points, queries=None, k: int = 1, max_radius: float = math.inf, threads: int = 0) -> np.ndarray:
    x = _f32c(points)
    q = x if queries is None else _f32c(queries)
    out = np.empty(len(q), np.float32)
    check(lib().b200knn_cpu_bruteforce(x.ctypes.data, len(x), q.ctypes.data, len(q), int(k),
                                       float(max_radius), out.ctypes.data, int(threads)), "cpu_bruteforce")
    return out


def cpu_knn(points, queries: Optional[np.ndarray] = None, k: int = 1, max_radius: float = math.inf,
            threads: int = 0) -> np.ndarray:
    x = _f32c(points)
    q = None if queries is None else _f32c(queries)
    nq = len(x) if q is None else len(q)
    out = np.empty(nq, np.float32)
    check(lib().b200knn_cpu_knn(x.ctypes.data, len(x), None if q is None else q.ctypes.data, nq, int(k),
                                float(max_radius), out.ctypes.data, int(threads)), "cpu_knn")
    return out


def neighbors_bruteforce(points, queries=None, k: int = 1, max_radius: float = math.inf, threads: int = 0):
    """Brute-force k nearest neighbours: (ids int32 [nq,k], distances float32 [nq,k]), rows sorted by
    (distance, id); -1 / +inf where fewer than k points are in range."""
    x = _f32c(points)
    q = x if queries is None else _f32c(queries)
    ids = np.empty((len(q), int(k)), np.int32)
    dist = np.empty((len(q), int(k)), np.float32)
    fn = lib().b200knn_cpu_neighbors
    fn.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_void_p, C.c_void_p, C.c_int]
    fn.restype = C.c_int
    check(fn(x.ctypes.data, len(x), q.ctypes.data, len(q), int(k), float(max_radius), ids.ctypes.data,
             dist.ctypes.data, int(threads)), "cpu_neighbors")
    return ids, dist


def bit_equal(a: np.ndarray, b: np.ndarray) -> bool:
    a = np.ascontiguousarray(a, np.float32)
    b = np.ascontiguousarray(b, np.float32)
    return a.shape == b.shape and bool(np.array_equal(a.view(np.uint32), b.view(np.uint32)))
