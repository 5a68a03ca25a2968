"""GPU: k nearest neighbours with ids (two-pass: k-th distance, then fixed-radius collection)
against the brute-force neighbour oracle. Distances must match bit for bit; ids may differ only
among points at exactly equal distance, so they are checked by recomputing their distances."""
import math

import numpy as np
import pytest

torch = pytest.importorskip("torch")

from mpi_cuda_largescaleknn_b200.utils import generators as gen
from mpi_cuda_largescaleknn_b200.utils.oracle import bit_equal, bruteforce, neighbors_bruteforce

pytestmark = pytest.mark.gpu


def _check(x, q, k, r=math.inf):
    from mpi_cuda_largescaleknn_b200.ops.knn import KnnIndex, neighbors
    dev = torch.device("cuda", 0)
    idx = KnnIndex(torch.from_numpy(x).to(dev))
    ids, dist = neighbors(idx, None if q is None else torch.from_numpy(q).to(dev), k, r)
    ids, dist = ids.cpu().numpy(), dist.cpu().numpy()
    idx.close()
    qq = x if q is None else q
    want_ids, want_dist = neighbors_bruteforce(x, qq, k=k, max_radius=r)
    assert bit_equal(dist, want_dist)
    assert np.array_equal(ids >= 0, want_ids >= 0)
    # every reported id really lies at the reported distance, and no id is reported twice
    for row in range(0, len(qq), max(1, len(qq) // 200)):
        have = ids[row] >= 0
        d = np.array([bruteforce(x[i][None, :], queries=qq[row][None, :], k=1)[0] for i in ids[row][have]], np.float32)
        assert bit_equal(d, dist[row][have])
        assert len(set(ids[row][have].tolist())) == int(have.sum())


@pytest.mark.parametrize("n,k", [(1, 1), (40, 8), (5000, 1), (5000, 32), (20_000, 100), (3000, 500)])
def test_self_neighbors(n, k):
    _check(gen.uniform(n, seed=n + k), None, k)


def test_external_queries_and_radius():
    x = gen.uniform(30_000, seed=5)
    q = gen.uniform(4001, seed=6) * 1.3 - 0.15
    _check(x, q, 16)
    _check(x, q, 40, r=0.04)


def test_fewer_points_than_k_and_duplicates():
    _check(gen.uniform(10, seed=1), None, 16)
    x = np.concatenate([gen.uniform(2000, seed=2), np.full((300, 3), 0.5, np.float32)])
    _check(x, None, 64)          # 300 coincident points: rows full of exact ties


def test_counting_verifier_on_a_large_cloud():
    """verify_sample: 64 sampled results of a 4 M-point run checked against all points by counting; a corrupted
    result must be flagged."""
    from mpi_cuda_largescaleknn_b200.ops.knn import knn, verify_sample
    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cuda").manual_seed(11)
    x = torch.rand((4_000_000, 3), device=dev, generator=g)
    res = knn(x, 50)
    assert verify_sample(x, res, 50, samples=64, seed=3) == {"queries": 64, "violations": 0}
    bad = res.clone()
    bad *= 1.01
    assert verify_sample(x, bad, 50, samples=64, seed=3)["violations"] == 64
