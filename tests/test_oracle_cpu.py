"""CPU-only: the mirror engine (same algorithm as the GPU kernels) must equal the brute-force
oracle bit for bit; the oracle itself is cross-checked against float64 numpy."""
import math

import numpy as np
import pytest

from mpi_cuda_largescaleknn_b200.utils import generators as gen
from mpi_cuda_largescaleknn_b200.utils.oracle import bit_equal, bruteforce, cpu_knn


@pytest.mark.parametrize("n,k", [(1, 1), (31, 1), (32, 5), (33, 5), (1000, 10), (1025, 32), (4000, 100), (20, 50)])
def test_mirror_engine_equals_oracle_uniform(n, k):
    x = gen.uniform(n, seed=n + k)
    assert bit_equal(cpu_knn(x, k=k), bruteforce(x, k=k))


def test_self_is_included_k1_is_zero():
    x = gen.uniform(500, seed=3)
    assert np.all(bruteforce(x, k=1) == 0.0)
    assert np.all(cpu_knn(x, k=1) == 0.0)


def test_fewer_than_k_points_gives_inf():
    x = gen.uniform(7, seed=1)
    assert np.all(np.isinf(bruteforce(x, k=8)))
    assert np.all(np.isinf(cpu_knn(x, k=8)))


@pytest.mark.parametrize("r", [0.01, 0.05, 0.3])
def test_max_radius(r):
    x = gen.uniform(3000, seed=9)
    a, b = bruteforce(x, k=16, max_radius=r), cpu_knn(x, k=16, max_radius=r)
    assert bit_equal(a, b)
    assert np.all(a <= np.float32(math.sqrt(np.float32(r) * np.float32(r))) + 0)


def test_duplicates_and_clusters():
    x = np.concatenate([gen.clustered_surface(3000, seed=2), np.zeros((40, 3), np.float32)])
    assert bit_equal(cpu_knn(x, k=20), bruteforce(x, k=20))


def test_external_queries():
    x = gen.uniform(2500, seed=4)
    q = gen.uniform(333, seed=5) * 1.5 - 0.25
    assert bit_equal(cpu_knn(x, queries=q, k=7), bruteforce(x, queries=q, k=7))


def test_oracle_close_to_float64():
    x = gen.uniform(400, seed=6)
    d = np.sqrt(((x[:, None, :].astype(np.float64) - x[None, :, :]) ** 2).sum(-1))
    ref = np.sort(d, axis=1)[:, 9]
    got = bruteforce(x, k=10)
    assert np.max(np.abs(got - ref) / ref) < 1e-6


def test_oracle_uses_canonical_fma_order():
    # one query/point pair where the canonical order differs from naive x,y,z summation
    rng = np.random.default_rng(0)
    found = False
    for _ in range(2000):
        p = rng.random(3).astype(np.float32)
        q = rng.random(3).astype(np.float32)
        d = (q - p).astype(np.float32)
        canon = np.float32(math.fma(float(d[2]), float(d[2]), float(np.float32(math.fma(float(d[0]), float(d[0]), float(np.float32(d[1] * d[1]))))))) if hasattr(math, "fma") else None
        naive = np.float32(np.float32(np.float32(d[0] * d[0]) + np.float32(d[1] * d[1])) + np.float32(d[2] * d[2]))
        got = bruteforce(p[None, :], queries=q[None, :], k=1)[0]
        if canon is not None:
            assert got == np.float32(math.sqrt(canon)) or np.float32(np.sqrt(canon)) == got
            if canon != naive:
                found = True
    if hasattr(math, "fma"):
        assert found


def test_neighbors_bruteforce_rows_are_the_k_smallest_sorted():
    from mpi_cuda_largescaleknn_b200.utils.oracle import neighbors_bruteforce
    x = gen.uniform(700, seed=41)
    q = gen.uniform(50, seed=42)
    ids, dist = neighbors_bruteforce(x, q, k=9)
    d_all = np.sqrt(((q[:, None, :].astype(np.float64) - x[None, :, :]) ** 2).sum(-1))
    want = np.sort(d_all, axis=1)[:, :9]
    assert np.allclose(dist, want, rtol=1e-5, atol=1e-7)
    assert np.all(np.diff(dist, axis=1) >= 0)
    assert np.allclose(d_all[np.arange(50)[:, None], ids], dist, rtol=1e-5, atol=1e-7)
    # the k-th column is the k-th-distance oracle, bit for bit
    assert bit_equal(dist[:, -1], bruteforce(x, queries=q, k=9))
    # fewer than k points / cut-off radius: padding
    ids2, dist2 = neighbors_bruteforce(x[:5], q[:3], k=8)
    assert np.all(ids2[:, 5:] == -1) and np.all(np.isinf(dist2[:, 5:])) and np.all(ids2[:, :5] >= 0)
    ids3, dist3 = neighbors_bruteforce(x, q, k=9, max_radius=0.05)
    assert np.all((dist3 < 0.05) == (ids3 >= 0))


def test_two_dimensional_clouds_are_exact():
    """[N,2] input (the reference library's float2 trees): padding z = 0 reproduces fmaf(dx,dx, dy*dy) exactly."""
    rng = np.random.default_rng(5)
    x = rng.random((600, 2), dtype=np.float32)
    d2 = np.empty((600, 600), np.float32)
    for i in range(600):
        dx, dy = x[i, 0] - x[:, 0], x[i, 1] - x[:, 1]
        d2[i] = (dx.astype(np.float64) * dx.astype(np.float64) + (dy * dy).astype(np.float64)).astype(np.float32)   # fmaf(dx,dx,dy*dy)
    want = np.sqrt(np.sort(d2, axis=1)[:, 6])
    assert bit_equal(bruteforce(x, k=7), want)
    assert bit_equal(cpu_knn(x, k=7), want)
