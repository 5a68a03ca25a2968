"""The allocation-free entry point, called through the C ABI exactly as a C caller would.

Some callers of libb200knn cannot let the library allocate device memory: the host program has
already given the whole GPU to its own arena, or it runs a loop that must never hit the device-wide
stall a cudaMalloc/cudaFree can cause, or its plug-in contract forbids allocator calls after start-up.
For them ``b200knn_knn_self_noalloc`` keeps all working memory in one fixed array inside the library.
This test checks that it returns exactly what the normal, allocating entry points return
(``b200knn_index_create`` + ``b200knn_knn_self``), twice in a row — the second call finds the first
call's leftovers in the shared array and must not be disturbed by them — that oversized inputs are
refused with a clear message, and that a call does not change the amount of free device memory.

The file name sorts last on purpose: with ``pytest -x`` every other test has run before this one.
It must pass on its own in a fresh process as well as after the rest of the suite.
"""
import ctypes as C
import math

import numpy as np
import pytest

torch = pytest.importorskip("torch")

from mpi_cuda_largescaleknn_b200.ops._lib import lib
from mpi_cuda_largescaleknn_b200.utils import generators as gen

pytestmark = pytest.mark.gpu


def _abi():
    L = lib()
    L.b200knn_knn_self_noalloc.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_void_p, C.c_void_p]
    L.b200knn_knn_self_noalloc.restype = C.c_int
    L.b200knn_noalloc_max_points.argtypes = []
    L.b200knn_noalloc_max_points.restype = C.c_int64
    L.b200knn_noalloc_scratch_bytes.argtypes = []
    L.b200knn_noalloc_scratch_bytes.restype = C.c_int64
    return L


def _reference_path(L, x_dev, k, r):
    """The existing entry points: index_create + knn_self (these allocate)."""
    n = x_dev.shape[0]
    out = torch.empty(n, dtype=torch.float32, device=x_dev.device)
    h = C.c_void_p()
    stream = torch.cuda.current_stream().cuda_stream
    assert L.b200knn_index_create(x_dev.data_ptr(), n, None, 0, stream, C.byref(h)) == 0, L.b200knn_last_error()
    assert L.b200knn_knn_self(h, k, r, out.data_ptr(), stream) == 0, L.b200knn_last_error()
    torch.cuda.synchronize()
    L.b200knn_index_destroy(h)
    return out.cpu().numpy()


def _noalloc(L, x_dev, k, r):
    out = torch.full((x_dev.shape[0],), -1.0, dtype=torch.float32, device=x_dev.device)
    stream = torch.cuda.current_stream().cuda_stream
    rc = L.b200knn_knn_self_noalloc(x_dev.data_ptr(), x_dev.shape[0], k, r, out.data_ptr(), stream)
    assert rc == 0, L.b200knn_last_error()
    torch.cuda.synchronize()
    return out.cpu().numpy()


def test_noalloc_matches_allocating_path_twice_in_a_row():
    L = _abi()
    dev = torch.device("cuda", 0)
    assert L.b200knn_noalloc_max_points() == 262144
    assert 0 < L.b200knn_noalloc_scratch_bytes() < 16 * 1024 * 1024
    # two different problems back to back: the second runs on top of the first one's leftovers
    cases = [(200_000, 100, math.inf, gen.uniform(200_000, seed=1)),
             (70_001, 16, 0.05, gen.clustered_surface(70_001, seed=2))]
    for n, k, r, x in cases:
        xd = torch.from_numpy(x).to(dev)
        want = _reference_path(L, xd, k, r)
        first = _noalloc(L, xd, k, r)
        second = _noalloc(L, xd, k, r)
        assert np.array_equal(first.view(np.uint32), want.view(np.uint32))
        assert np.array_equal(second.view(np.uint32), want.view(np.uint32))
    # a small problem after a large one: stale boxes / points beyond n must not leak in
    x = gen.uniform(33, seed=3)
    xd = torch.from_numpy(x).to(dev)
    assert np.array_equal(_noalloc(L, xd, 5, math.inf).view(np.uint32),
                          _reference_path(L, xd, 5, math.inf).view(np.uint32))


def test_noalloc_rejects_oversized_input_and_does_not_allocate():
    L = _abi()
    dev = torch.device("cuda", 0)
    big = torch.zeros((262145, 3), dtype=torch.float32, device=dev)
    out = torch.zeros(262145, dtype=torch.float32, device=dev)
    rc = L.b200knn_knn_self_noalloc(big.data_ptr(), 262145, 4, math.inf, out.data_ptr(), None)
    assert rc != 0
    assert b"B200KNN_NOALLOC_MAX_POINTS" in L.b200knn_last_error()
    # k beyond the shared-memory lists would need the allocating fallback: refused, and before anything runs
    small = torch.from_numpy(gen.uniform(5_000, seed=5)).to(dev)
    out_small = torch.zeros(5_000, dtype=torch.float32, device=dev)
    rc = L.b200knn_knn_self_noalloc(small.data_ptr(), 5_000, 2000, math.inf, out_small.data_ptr(), None)
    assert rc != 0 and b"does not fit" in L.b200knn_last_error()
    # warm call (first use may load kernels), then neither the driver's free memory nor the library's
    # stream-ordered pool may move across a call: the pool keeps freed blocks, so a cudaMallocAsync would
    # not show in mem_get_info — it shows in the pool's high-water mark of bytes in use
    L.b200knn_mempool_usage.argtypes = [C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.c_int]
    L.b200knn_mempool_usage.restype = C.c_int
    x = torch.from_numpy(gen.uniform(100_000, seed=4)).to(dev)
    _noalloc(L, x, 10, math.inf)
    torch.cuda.synchronize()
    reserved0, high0 = C.c_int64(), C.c_int64()
    assert L.b200knn_mempool_usage(C.byref(reserved0), C.byref(high0), 1) == 0
    free_before, _ = torch.cuda.mem_get_info(dev)
    _noalloc(L, x, 10, math.inf)
    rc = L.b200knn_knn_self_noalloc(small.data_ptr(), 5_000, 2000, math.inf, out_small.data_ptr(), None)
    assert rc != 0
    free_after, _ = torch.cuda.mem_get_info(dev)
    reserved1, high1 = C.c_int64(), C.c_int64()
    assert L.b200knn_mempool_usage(C.byref(reserved1), C.byref(high1), 0) == 0
    assert free_after == free_before
    assert reserved1.value == reserved0.value
    assert high1.value == 0, "the call used the stream-ordered pool"
