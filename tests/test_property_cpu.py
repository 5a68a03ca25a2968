"""CPU-only property tests: the mirror engine (same index, same group traversal, same lane-set
splitting as the GPU kernels) must agree bit for bit with the brute-force oracle on whatever small
cloud hypothesis comes up with — duplicates, collinear and coplanar sets, huge and tiny coordinates,
cut-off radii, k above n."""
import math

import numpy as np
import pytest

hyp = pytest.importorskip("hypothesis")
from hypothesis import given, settings  # noqa: E402
from hypothesis import strategies as st  # noqa: E402

from mpi_cuda_largescaleknn_b200.utils.oracle import bit_equal, bruteforce, cpu_knn  # noqa: E402

finite = st.floats(min_value=-1e6, max_value=1e6, allow_nan=False, allow_infinity=False, width=32)


@st.composite
def clouds(draw):
    n = draw(st.integers(min_value=1, max_value=400))
    kind = draw(st.sampled_from(["random", "grid", "line", "plane", "two_clusters", "all_equal"]))
    seed = draw(st.integers(min_value=0, max_value=2**31 - 1))
    rng = np.random.default_rng(seed)
    scale = np.float32(draw(st.sampled_from([1e-3, 1.0, 1e3, 1e6])))
    if kind == "random":
        x = rng.random((n, 3), dtype=np.float32)
    elif kind == "grid":                       # many exact ties
        x = rng.integers(0, 4, (n, 3)).astype(np.float32)
    elif kind == "line":
        x = np.zeros((n, 3), np.float32)
        x[:, 0] = rng.random(n, dtype=np.float32)
    elif kind == "plane":
        x = rng.random((n, 3), dtype=np.float32)
        x[:, 2] = np.float32(0.25)
    elif kind == "two_clusters":               # the case that makes query groups straddle a gap
        x = rng.normal(0, 0.01, (n, 3)).astype(np.float32)
        x[n // 2:] += np.float32(50.0)
    else:
        x = np.full((n, 3), draw(finite), np.float32)
    return np.ascontiguousarray(x * scale)


@settings(max_examples=80, deadline=None)
@given(clouds(), st.integers(min_value=1, max_value=70), st.sampled_from([math.inf, 1e-4, 0.3, 10.0]))
def test_mirror_engine_self_queries_equal_oracle(x, k, r):
    assert bit_equal(cpu_knn(x, k=k, max_radius=r, threads=1), bruteforce(x, k=k, max_radius=r, threads=1))


@settings(max_examples=40, deadline=None)
@given(clouds(), clouds(), st.integers(min_value=1, max_value=40))
def test_mirror_engine_external_queries_equal_oracle(x, q, k):
    assert bit_equal(cpu_knn(x, queries=q, k=k, threads=1), bruteforce(x, queries=q, k=k, threads=1))
