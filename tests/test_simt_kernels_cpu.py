"""CPU-only: the CUDA kernels themselves, run under the host SIMT emulation (tests/simt): the same
.cu sources compiled as C++, every CUDA thread a fiber, warp collectives implemented as exchanges
between the 32 fibers of a warp. Checks the kernels' logic — radix sort, k-d refine, box tree, the
group traversal with its heaps, frontier and fill phase, the culled-exchange and redistribution
kernels on one rank, the no-allocation entry point — bit for bit against the brute-force oracle,
on the machine that has no GPU. (Memory-model behaviour and speed are what only the GPU tests see.)"""
import ctypes as C
import math
import platform
import subprocess
from pathlib import Path

import numpy as np
import pytest

from conftest import run_make

from mpi_cuda_largescaleknn_b200.ops import _lib
from mpi_cuda_largescaleknn_b200.utils import generators as gen
from mpi_cuda_largescaleknn_b200.utils.oracle import bit_equal, bruteforce

ROOT = Path(__file__).resolve().parents[1]
pytestmark = pytest.mark.skipif(platform.machine() != "x86_64", reason="the emulation's fiber switch is x86-64 assembly")


@pytest.fixture(scope="module")
def sim():
    out = run_make("simt")
    assert out.returncode == 0, out.stdout[-3000:]
    import os
    lib = _lib._proto(C.CDLL(os.environ.get("B200KNN_SIMT_LIB", str(ROOT / "build" / "libb200knn_simt.so"))))
    lib.b200knn_knn_self_noalloc.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_void_p, C.c_void_p]
    lib.b200knn_knn_self_noalloc.restype = C.c_int
    return lib


def ok(lib, rc):
    assert rc == 0, lib.b200knn_last_error().decode()


class Index:
    def __init__(self, lib, x):
        self.lib, self.x, self.h = lib, np.ascontiguousarray(x, np.float32), C.c_void_p()
        ok(lib, lib.b200knn_index_create(self.x.ctypes.data, len(self.x), None, 0, None, C.byref(self.h)))

    def knn_self(self, k, r=math.inf):
        out = np.full(len(self.x), -1.0, np.float32)
        ok(self.lib, self.lib.b200knn_knn_self(self.h, k, r, out.ctypes.data, None))
        return out

    def knn_query(self, q, k, r=math.inf):
        q = np.ascontiguousarray(q, np.float32)
        out = np.full(len(q), -1.0, np.float32)
        ok(self.lib, self.lib.b200knn_knn_query(self.h, q.ctypes.data, len(q), k, r, out.ctypes.data, None))
        return out

    def close(self):
        self.lib.b200knn_index_destroy(self.h)


@pytest.mark.parametrize("n,k", [(1, 1), (31, 3), (33, 8), (20, 50), (1000, 10), (2500, 32), (3000, 100), (1500, 221)])
def test_emulated_self_knn_equals_oracle(sim, n, k):
    x = gen.uniform(n, seed=7 * n + k)
    idx = Index(sim, x)
    assert bit_equal(idx.knn_self(k), bruteforce(x, k=k))
    idx.close()


def test_emulated_index_is_sorted_and_boxes_are_exact(sim):
    n = 5000
    x = gen.uniform(n, seed=3)
    idx = Index(sim, x)
    s = np.empty((n, 4), np.float32)
    ok(sim, sim.b200knn_index_copy_sorted(idx.h, s.ctypes.data))
    ids = s[:, 3].copy().view(np.uint32)
    assert np.array_equal(np.sort(ids), np.arange(n, dtype=np.uint32))
    assert np.array_equal(s[:, :3], x[ids])
    nl = (n + 31) // 32
    boxes = np.empty((nl, 8), np.float32)
    cnt = C.c_int()
    ok(sim, sim.b200knn_index_copy_boxes(idx.h, 0, boxes.ctypes.data, C.byref(cnt)))
    assert cnt.value == nl
    pts = np.concatenate([s[:, :3], np.full((nl * 32 - n, 3), np.nan, np.float32)]).reshape(nl, 32, 3)
    assert np.array_equal(boxes[:, 0:3], np.nanmin(pts, axis=1)) and np.array_equal(boxes[:, 4:7], np.nanmax(pts, axis=1))
    # k-d refinement: every full 1024-point run is 5 levels of exact median splits along the widest axis

    def check_kd(seg, depth):
        if depth == 5:
            return
        axis = int(np.argmax(seg.max(0) - seg.min(0)))
        half = len(seg) // 2
        assert seg[:half, axis].max() <= seg[half:, axis].min()
        check_kd(seg[:half], depth + 1)
        check_kd(seg[half:], depth + 1)

    for run in range(n // 1024):
        check_kd(s[run * 1024:(run + 1) * 1024, :3], 0)
    idx.close()


def test_emulated_index_sort_falls_back_to_full_keys_on_heavy_cells(sim):
    """The build sorts by the leading 40 bits of the Morton key and, when more than 1024 points share one such cell,
    again by all 64: exact results either way, and the switch must happen exactly when it should."""
    sim.b200knn_index_sort_bits.argtypes = [C.c_void_p]
    x = gen.uniform(3000, seed=5)
    tight = (np.float32(0.5) + np.random.default_rng(6).random((1500, 3), dtype=np.float32) * np.float32(1e-6)).astype(np.float32)
    both = np.ascontiguousarray(np.concatenate([x, tight]))
    a, b = Index(sim, x), Index(sim, both)
    assert sim.b200knn_index_sort_bits(a.h) == 40 and sim.b200knn_index_sort_bits(b.h) == 64
    assert bit_equal(b.knn_self(12), bruteforce(both, k=12))
    few = np.ascontiguousarray(np.concatenate([x, tight[:900]]))      # 900 in the cell: no need
    c = Index(sim, few)
    assert sim.b200knn_index_sort_bits(c.h) == 40
    assert bit_equal(c.knn_self(12), bruteforce(few, k=12))
    for i in (a, b, c):
        i.close()


def test_emulated_radius_duplicates_and_clusters(sim):
    x = np.concatenate([gen.clustered_surface(2000, seed=2), np.zeros((60, 3), np.float32),
                        np.full((40, 3), 0.25, np.float32)])
    idx = Index(sim, x)
    assert bit_equal(idx.knn_self(32), bruteforce(x, k=32))
    assert bit_equal(idx.knn_self(40, 0.05), bruteforce(x, k=40, max_radius=0.05))      # fill path with a cut-off
    idx.close()
    # two far-apart clusters: query groups straddle the gap and are split before traversal
    y = np.concatenate([gen.uniform(700, seed=4) * 0.01, gen.uniform(700, seed=5) * 0.01 + 50.0]).astype(np.float32)
    np.random.default_rng(0).shuffle(y)
    idx = Index(sim, y)
    assert bit_equal(idx.knn_self(20), bruteforce(y, k=20))
    idx.close()


def test_emulated_external_queries_and_fcp(sim):
    x = gen.uniform(3000, seed=8)
    q = gen.uniform(701, seed=9) * 1.4 - 0.2
    idx = Index(sim, x)
    assert bit_equal(idx.knn_query(q, 12), bruteforce(x, queries=q, k=12))
    assert bit_equal(idx.knn_query(q, 40), bruteforce(x, queries=q, k=40))
    ids = np.empty(len(q), np.int32)
    dist = np.empty(len(q), np.float32)
    ok(sim, sim.b200knn_fcp_query(idx.h, q.ctypes.data, len(q), math.inf, ids.ctypes.data, dist.ctypes.data, None))
    assert bit_equal(dist, bruteforce(x, queries=q, k=1))
    assert np.allclose(np.sqrt(((x[ids] - q) ** 2).sum(1)), dist, rtol=1e-5)
    idx.close()


def test_emulated_index_over_records_with_payload(sim):
    """b200knn_index_create_records: positions read in place from records with payload around them; results equal
    the oracle on the positions, ids are record numbers, bad layouts are refused."""
    rng = np.random.default_rng(3)
    for dims, first, width in ((3, 2, 7), (2, 1, 4)):
        rec = rng.random((2600, width), dtype=np.float32)
        pos = np.zeros((len(rec), 3), np.float32)
        pos[:, :dims] = rec[:, first:first + dims]
        q = rng.random((300, dims), dtype=np.float32)
        q3 = np.zeros((len(q), 3), np.float32)
        q3[:, :dims] = q
        h = C.c_void_p()
        sim.b200knn_index_create_records.argtypes = [C.c_void_p, C.c_int64, C.c_size_t, C.c_size_t, C.c_int, C.c_void_p,
                                                     C.c_uint32, C.c_void_p, C.POINTER(C.c_void_p)]
        ok(sim, sim.b200knn_index_create_records(rec.ctypes.data, len(rec), width * 4, first * 4, dims, None, 0, None, C.byref(h)))
        out = np.empty(len(rec), np.float32)
        ok(sim, sim.b200knn_knn_self(h, 20, math.inf, out.ctypes.data, None))
        assert bit_equal(out, bruteforce(pos, k=20))
        ids, dist = np.empty(len(q), np.int32), np.empty(len(q), np.float32)
        ok(sim, sim.b200knn_fcp_query(h, q.ctypes.data, len(q), math.inf, ids.ctypes.data, dist.ctypes.data, None))
        assert bit_equal(dist, bruteforce(pos, queries=q3, k=1))
        assert np.allclose(np.sqrt(((rec[ids][:, first:first + dims] - q) ** 2).sum(1)), dist, rtol=1e-5, atol=1e-7)
        sim.b200knn_index_destroy(h)
    h = C.c_void_p()
    assert sim.b200knn_index_create_records(rec.ctypes.data, len(rec), 16, 8, 3, None, 0, None, C.byref(h)) != 0   # 8 + 12 > 16
    assert sim.b200knn_index_create_records(rec.ctypes.data, len(rec), 14, 0, 3, None, 0, None, C.byref(h)) != 0   # not a multiple of 4
    assert sim.b200knn_index_create_records(rec.ctypes.data, len(rec), 16, 0, 4, None, 0, None, C.byref(h)) != 0   # dims


def test_emulated_heap_in_global_memory(sim):
    x = gen.uniform(2300, seed=77)
    idx = Index(sim, x)
    assert bit_equal(idx.knn_self(1900), bruteforce(x, k=1900))
    idx.close()


def test_emulated_no_alloc_entry_point(sim):
    x = gen.uniform(4000, seed=11)
    out = np.empty(len(x), np.float32)
    for _ in range(2):
        ok(sim, sim.b200knn_knn_self_noalloc(x.ctypes.data, len(x), 16, math.inf, out.ctypes.data, None))
        assert bit_equal(out, bruteforce(x, k=16))


def test_emulated_distributed_driver_single_rank(sim):
    """world = 1: culled-exchange kernels (LOCAL phase), redistribution scatter, result return."""
    from mpi_cuda_largescaleknn_b200.parallel.dist import Bootstrap, ALLGATHER_FN, BARRIER_FN, DistReport, MODES
    x = gen.uniform(6000, seed=31)
    bs = Bootstrap(None, 0, 1, ALLGATHER_FN(lambda *a: 1), BARRIER_FN(lambda *a: 1))
    eng = C.c_void_p()
    sim.b200knn_dist_create.argtypes = [C.POINTER(Bootstrap), C.POINTER(C.c_void_p)]
    ok(sim, sim.b200knn_dist_create(C.byref(bs), C.byref(eng)))
    sim.b200knn_dist_knn.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int, C.c_float, C.c_int, C.c_int,
                                     C.c_void_p, C.c_void_p, C.POINTER(DistReport)]
    want = bruteforce(x, k=40)
    for mode in ("prepartitioned", "redistribute", "auto"):
        out = np.empty(len(x), np.float32)
        rep = DistReport()
        ok(sim, sim.b200knn_dist_knn(eng, x.ctypes.data, len(x), 0, 40, math.inf, MODES[mode], 0, out.ctypes.data, None,
                                     C.byref(rep)))
        assert bit_equal(out, want), mode
    sim.b200knn_dist_destroy.argtypes = [C.c_void_p]
    sim.b200knn_dist_destroy(eng)


def _neighbors(lib, idx, q, k, r=math.inf):
    fn = lib.b200knn_knn_neighbors
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
    fn.restype = C.c_int
    nq = len(idx.x) if q is None else len(q)
    ids = np.full((nq, k), -7, np.int32)
    dist = np.full((nq, k), -7.0, np.float32)
    ok(lib, fn(idx.h, None if q is None else q.ctypes.data, nq, k, r, ids.ctypes.data, dist.ctypes.data, None))
    order = np.argsort(dist, axis=1, kind="stable")
    return np.take_along_axis(ids, order, 1), np.take_along_axis(dist, order, 1)


@pytest.mark.parametrize("n,k,r", [(1, 1, math.inf), (40, 8, math.inf), (10, 16, math.inf), (3000, 1, math.inf),
                                   (3000, 32, math.inf), (2500, 100, math.inf), (3000, 40, 0.05)])
def test_emulated_neighbor_lists(sim, n, k, r):
    """k nearest neighbours WITH ids: distances bit-identical to the brute-force neighbour oracle, every id lies
    at its reported distance, no id twice, padding where fewer than k points are in range."""
    from mpi_cuda_largescaleknn_b200.utils.oracle import neighbors_bruteforce
    x = gen.uniform(n, seed=n + k)
    idx = Index(sim, x)
    for q in (None, (gen.uniform(333, seed=k) * 1.3 - 0.15).astype(np.float32)):
        ids, dist = _neighbors(sim, idx, q, k, r)
        qq = x if q is None else q
        want_ids, want_dist = neighbors_bruteforce(x, qq, k=k, max_radius=r)
        assert bit_equal(dist, want_dist)
        assert np.array_equal(ids >= 0, want_ids >= 0)
        for row in range(0, len(qq), max(1, len(qq) // 150)):
            have = ids[row] >= 0
            d = np.array([bruteforce(x[i][None, :], queries=qq[row][None, :], k=1)[0] for i in ids[row][have]], np.float32)
            assert bit_equal(d, dist[row][have])
            assert len(set(ids[row][have].tolist())) == int(have.sum())
    idx.close()


def test_emulated_neighbor_lists_with_exact_ties(sim):
    from mpi_cuda_largescaleknn_b200.utils.oracle import neighbors_bruteforce
    x = np.concatenate([gen.uniform(800, seed=2), np.full((300, 3), 0.5, np.float32),
                        np.random.default_rng(1).integers(0, 3, (400, 3)).astype(np.float32)])
    idx = Index(sim, x)
    ids, dist = _neighbors(sim, idx, None, 64)
    want_ids, want_dist = neighbors_bruteforce(x, x, k=64)
    assert bit_equal(dist, want_dist)
    for row in range(0, len(x), 11):
        assert len(set(ids[row].tolist())) == 64
        d = np.array([bruteforce(x[i][None, :], queries=x[row][None, :], k=1)[0] for i in ids[row]], np.float32)
        assert bit_equal(d, dist[row])
    idx.close()


def test_emulated_frontier_overflow_paths_are_exercised_and_exact():
    """The frontier (inner nodes only since leaves stay in registers) never overflows its 128 entries on small
    inputs, so the test shrinks it to 8 (B200KNN_QUEUE_CAP, read once per process: hence the child process):
    k = 200 on 40 000 points then purges stale entries and falls back to the depth-first frame stack. The
    counters prove the paths ran; the result stays exact."""
    import os, sys
    code = (
        "import ctypes as C, numpy as np, os\n"
        "from mpi_cuda_largescaleknn_b200.ops import _lib\n"
        "from mpi_cuda_largescaleknn_b200.utils import generators as gen\n"
        "from mpi_cuda_largescaleknn_b200.utils.oracle import bit_equal, bruteforce\n"
        f"lib = _lib._proto(C.CDLL(os.environ.get('B200KNN_SIMT_LIB', r'{ROOT / 'build' / 'libb200knn_simt.so'}')))\n"
        "x = gen.uniform(40_000, seed=1)\n"
        "h = C.c_void_p()\n"
        "assert lib.b200knn_index_create(x.ctypes.data, len(x), None, 0, None, C.byref(h)) == 0\n"
        "assert lib.b200knn_index_enable_stats(h, 1) == 0\n"
        "out = np.empty(len(x), np.float32)\n"
        "assert lib.b200knn_knn_self(h, 200, float('inf'), out.ctypes.data, None) == 0\n"
        "st = (C.c_uint64 * 8)()\n"
        "assert lib.b200knn_index_stats(h, st) == 0\n"
        "assert st[4] > 0 and st[5] > 0, f'fallbacks {st[4]}, purges {st[5]}'\n"
        "sel = np.arange(0, len(x), 150)\n"
        "assert bit_equal(out[sel], bruteforce(x, queries=x[sel], k=200))\n"
        "print('OVERFLOW_OK', st[4], st[5])\n")
    out = run_make("simt")
    assert out.returncode == 0, out.stdout[-3000:]
    run = subprocess.run([sys.executable, "-c", code], cwd=str(ROOT), env=dict(os.environ, B200KNN_QUEUE_CAP="8"),
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1500)
    assert run.returncode == 0 and "OVERFLOW_OK" in run.stdout, run.stdout[-3000:]


def test_emulated_kernels_on_degenerate_clouds(sim):
    """Property-style sweep through the real kernel code: lines, planes, grids full of exact ties, two far
    clusters, everything identical, huge and tiny coordinate scales."""
    rng = np.random.default_rng(123)
    for trial in range(12):
        n = int(rng.integers(1, 900))
        kind = ["random", "grid", "line", "plane", "two_clusters", "all_equal"][trial % 6]
        if kind == "random":
            x = rng.random((n, 3), dtype=np.float32)
        elif kind == "grid":
            x = rng.integers(0, 4, (n, 3)).astype(np.float32)
        elif kind == "line":
            x = np.zeros((n, 3), np.float32)
            x[:, 0] = rng.random(n, dtype=np.float32)
        elif kind == "plane":
            x = rng.random((n, 3), dtype=np.float32)
            x[:, 2] = np.float32(0.25)
        elif kind == "two_clusters":
            x = rng.normal(0, 0.01, (n, 3)).astype(np.float32)
            x[n // 2:] += np.float32(50.0)
        else:
            x = np.full((n, 3), np.float32(rng.normal()), np.float32)
        x = np.ascontiguousarray(x * np.float32([1e-3, 1.0, 1e3, 1e6][trial % 4]))
        k = int(rng.integers(1, 70))
        r = [math.inf, 0.3, math.inf, 10.0][trial % 4]
        idx = Index(sim, x)
        assert bit_equal(idx.knn_self(k, r), bruteforce(x, k=k, max_radius=r)), (trial, kind, n, k, r)
        idx.close()


def test_emulated_counting_verifier(sim):
    """count_within proves a k-th distance right (less < k <= less_equal) and a wrong one wrong."""
    x = np.ascontiguousarray(np.concatenate([gen.uniform(5000, seed=3), np.full((40, 3), 0.5, np.float32)]))
    k = 25
    pick = np.arange(0, len(x), 101)[:64]
    q = np.ascontiguousarray(x[pick])
    r = bruteforce(x, queries=q, k=k)
    fn = sim.b200knn_count_within
    fn.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    fn.restype = C.c_int

    def counts(radius):
        radius = np.ascontiguousarray(radius, np.float32)
        less, le = np.full(len(q), 7, np.uint64), np.full(len(q), 7, np.uint64)
        ok(sim, fn(x.ctypes.data, len(x), q.ctypes.data, len(q), radius.ctypes.data, less.ctypes.data, le.ctypes.data, None))
        return less.astype(np.int64), le.astype(np.int64)

    less, le = counts(r)
    assert np.all(less < k) and np.all(le >= k)
    # one float below the true answer: fewer than k points are within it (unless the answer is 0)
    less2, le2 = counts(np.nextafter(r, np.float32(-1)))
    assert np.all((le2 < k) | (r == 0))
    # one float above: at least k points are strictly closer
    less3, _ = counts(np.nextafter(r, np.float32(np.inf)))
    assert np.all(less3 >= k)
    # split across "ranks": the counts add up
    la, ea = counts(r)
    half = len(x) // 2
    l1, e1 = np.zeros(len(q), np.uint64), np.zeros(len(q), np.uint64)
    l2, e2 = np.zeros(len(q), np.uint64), np.zeros(len(q), np.uint64)
    xa, xb = np.ascontiguousarray(x[:half]), np.ascontiguousarray(x[half:])
    ok(sim, fn(xa.ctypes.data, len(xa), q.ctypes.data, len(q), r.ctypes.data, l1.ctypes.data, e1.ctypes.data, None))
    ok(sim, fn(xb.ctypes.data, len(xb), q.ctypes.data, len(q), r.ctypes.data, l2.ctypes.data, e2.ctypes.data, None))
    assert np.array_equal((l1 + l2).astype(np.int64), la) and np.array_equal((e1 + e2).astype(np.int64), ea)
