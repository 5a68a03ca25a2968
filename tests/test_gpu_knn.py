"""GPU numerics: every CUDA path is compared bit-for-bit with the CPU brute-force oracle
(plain fp32 reference of the same op)."""
import math

import numpy as np
import pytest

torch = pytest.importorskip("torch")

from mpi_cuda_largescaleknn_b200.utils import generators as gen
from mpi_cuda_largescaleknn_b200.utils.oracle import bit_equal, bruteforce

pytestmark = pytest.mark.gpu


def _dev():
    return torch.device("cuda", 0)


def _knn(x, k, r=math.inf):
    from mpi_cuda_largescaleknn_b200.ops.knn import knn
    return knn(torch.from_numpy(x).to(_dev()), k, r).cpu().numpy()


def test_library_is_native_and_loaded():
    from mpi_cuda_largescaleknn_b200.ops._lib import LIB_PATH, lib
    assert lib().b200knn_version() >= 100
    maps = open("/proc/self/maps").read()
    assert str(LIB_PATH) in maps


def morton63_np(x, lo, hi):
    ext = hi - lo
    inv = np.where(ext > 0, np.float32(1.0) / ext, np.float32(0)).astype(np.float32)
    t = np.clip((x - lo) * inv, 0, 1).astype(np.float32)
    q = np.minimum((t * np.float32(2097151.0)).astype(np.uint64), np.uint64(2097151))
    def spread(v):
        v = v & np.uint64(0x1fffff)
        v = (v | (v << np.uint64(32))) & np.uint64(0x1f00000000ffff)
        v = (v | (v << np.uint64(16))) & np.uint64(0x1f0000ff0000ff)
        v = (v | (v << np.uint64(8))) & np.uint64(0x100f00f00f00f00f)
        v = (v | (v << np.uint64(4))) & np.uint64(0x10c30c30c30c30c3)
        v = (v | (v << np.uint64(2))) & np.uint64(0x1249249249249249)
        return v
    return (spread(q[:, 0]) << np.uint64(2)) | (spread(q[:, 1]) << np.uint64(1)) | spread(q[:, 2])


@pytest.mark.parametrize("n", [1, 31, 32, 33, 4095, 4096, 4097, 100_003, 1_000_000])
def test_index_build_sort_and_boxes(n):
    """Radix sort + gather + box hierarchy: sorted by Morton key, a permutation of the input,
    every box is the exact union of its children (the reference's checkTree idea,
    cudaKDTree/testing/floatN-knn-and-fcp.cu:338-377)."""
    from mpi_cuda_largescaleknn_b200.ops.knn import KnnIndex
    x = gen.uniform(n, seed=n)
    idx = KnnIndex(torch.from_numpy(x).to(_dev()))
    b = idx.bounds()
    assert np.array_equal(b[:3], x.min(0)) and np.array_equal(b[3:], x.max(0))
    s = idx.sorted_points()
    ids = s[:, 3].copy().view(np.uint32)
    assert np.array_equal(np.sort(ids), np.arange(n, dtype=np.uint32))
    assert np.array_equal(s[:, :3], x[ids])
    # Order: Morton across 1024-point runs (radix sort on the key's leading sort_bits), k-d median order inside a
    # run (refine).
    shift = np.uint64(64 - idx.sort_bits)
    assert idx.sort_bits in (40, 64)
    keys = morton63_np(s[:, :3], b[:3], b[3:]) >> shift
    nruns = (n + 1023) // 1024
    run_min = np.array([keys[r * 1024:(r + 1) * 1024].min() for r in range(nruns)])
    run_max = np.array([keys[r * 1024:(r + 1) * 1024].max() for r in range(nruns)])
    assert np.all(run_min[1:] >= run_max[:-1])
    # every run holds exactly the points a stable Morton sort of the input puts there
    ref_order = np.argsort(morton63_np(x, b[:3], b[3:]) >> shift, kind="stable").astype(np.uint32)
    for r in range(0, nruns, max(1, nruns // 64)):
        assert np.array_equal(np.sort(ids[r * 1024:(r + 1) * 1024]), np.sort(ref_order[r * 1024:(r + 1) * 1024]))
    # inside a full run: first median split is exact (left half <= right half on the widest axis)
    if n >= 1024:
        run = s[:1024, :3]
        axis = int(np.argmax(run.max(0) - run.min(0)))
        assert run[:512, axis].max() <= run[512:, axis].min()
    # leaf boxes
    boxes = idx.boxes(0)
    nl = (n + 31) // 32
    assert len(boxes) == nl
    pad = np.full((nl * 32 - n, 3), np.nan, np.float32)
    pts = np.concatenate([s[:, :3], pad]).reshape(nl, 32, 3)
    assert np.array_equal(boxes[:, 0:3], np.nanmin(pts, axis=1))
    assert np.array_equal(boxes[:, 4:7], np.nanmax(pts, axis=1))
    cnt = boxes[:, 3].copy().view(np.int32)
    assert cnt.sum() == n and np.all(cnt[:-1] == 32)
    prev = boxes
    for lvl in range(1, idx.num_levels):
        cur = idx.boxes(lvl)
        assert len(cur) == (len(prev) + 31) // 32
        for p in range(len(cur)):
            ch = prev[p * 32:(p + 1) * 32]
            assert np.array_equal(cur[p, 0:3], ch[:, 0:3].min(0))
            assert np.array_equal(cur[p, 4:7], ch[:, 4:7].max(0))
        prev = cur
    assert len(prev) <= 32
    idx.close()


@pytest.mark.parametrize("n,k", [(1, 1), (31, 3), (32, 1), (33, 8), (1000, 10), (5000, 100), (20, 50),
                                 (4097, 7), (30_000, 32), (100_000, 100)])
def test_knn_self_equals_oracle(n, k):
    x = gen.uniform(n, seed=11 * n + k)
    got = _knn(x, k)
    if n <= 30_000:
        assert bit_equal(got, bruteforce(x, k=k))
    else:
        sel = np.random.default_rng(0).choice(n, 2000, replace=False)
        assert bit_equal(got[sel], bruteforce(x, queries=x[sel], k=k))


def test_knn_k1_is_zero_and_small_n_inf():
    x = gen.uniform(3000, seed=5)
    assert np.all(_knn(x, 1) == 0.0)
    y = gen.uniform(9, seed=6)
    assert np.all(np.isinf(_knn(y, 10)))


@pytest.mark.parametrize("r", [0.01, 0.05, 0.3])
def test_knn_max_radius(r):
    x = gen.uniform(20_000, seed=7)
    assert bit_equal(_knn(x, 16, r), bruteforce(x, k=16, max_radius=r))


def test_knn_max_radius_with_k_spanning_several_leaves():
    """k >= 32 takes the convergent-fill path when it is enabled: values beyond the cut-off must be
    clamped to r*r exactly as a sifted insert would have refused them."""
    x = gen.uniform(20_000, seed=17)
    for k, r in ((64, 0.05), (100, 0.02), (40, 0.3)):
        assert bit_equal(_knn(x, k, r), bruteforce(x, k=k, max_radius=r))


def test_dist_engine_world1_every_mode_equals_oracle():
    """The distributed driver with a single rank (no process group): exercises the culled-exchange
    kernels' LOCAL phase and the redistribution pipeline on one GPU."""
    from mpi_cuda_largescaleknn_b200.parallel.dist import DistKnn
    x = gen.uniform(60_000, seed=31)
    sel = np.arange(0, len(x), 37)
    ref = bruteforce(x, queries=x[sel], k=40)
    eng = DistKnn()
    for mode in ("ring", "prepartitioned", "redistribute", "auto"):
        got = eng.knn(torch.from_numpy(x).to(_dev()), 40, mode=mode).cpu().numpy()
        assert bit_equal(got[sel], ref), mode
    eng.close()


def test_n_times_k_beyond_int32():
    """n * k = 2.2e9 > 2^31: the reference overflows its int offsets here
    (/root/reference/unorderedDataVariant.cu:80); every offset in this library is 64-bit."""
    n, k = 22_000_000, 100
    g = torch.Generator(device="cuda").manual_seed(3)
    x = torch.rand((n, 3), device=_dev(), generator=g)
    from mpi_cuda_largescaleknn_b200.ops.knn import knn
    got = knn(x, k)
    sel = torch.arange(0, n, n // 96, device=_dev())
    xh = x.cpu().numpy()
    # 16 oracle threads: each holds all n distances of one query (88 MB)
    assert bit_equal(got[sel].cpu().numpy(), bruteforce(xh, queries=xh[sel.cpu().numpy()], k=k, threads=16))


def test_knn_clustered_duplicates():
    x = np.concatenate([gen.clustered_surface(20_000, seed=2), np.zeros((100, 3), np.float32),
                        np.full((70, 3), 0.25, np.float32)])
    assert bit_equal(_knn(x, 32), bruteforce(x, k=32))


def test_knn_negative_and_large_coordinates():
    rng = np.random.default_rng(3)
    x = (rng.normal(0, 1000.0, (8000, 3))).astype(np.float32)
    assert bit_equal(_knn(x, 10), bruteforce(x, k=10))


def test_knn_degenerate_plane_and_line():
    rng = np.random.default_rng(4)
    x = rng.random((5000, 3), dtype=np.float32)
    x[:, 2] = 0.5
    assert bit_equal(_knn(x, 9), bruteforce(x, k=9))
    x[:, 1] = -1.0
    assert bit_equal(_knn(x, 9), bruteforce(x, k=9))


@pytest.mark.parametrize("k", [221, 500])
def test_knn_large_k_uses_fewer_warps(k):
    x = gen.uniform(3000, seed=k)
    assert bit_equal(_knn(x, k), bruteforce(x, k=k))


def test_knn_k_beyond_shared_memory_uses_global_heaps():
    """k > ~1790 does not fit the shared-memory heap: the kernel variant with heaps in global memory
    must give the same bits (the reference supports any k that fits device memory)."""
    x = gen.uniform(6000, seed=77)
    assert bit_equal(_knn(x, 2000), bruteforce(x, k=2000))


def test_knn_query_external_points():
    from mpi_cuda_largescaleknn_b200.ops.knn import KnnIndex
    x = gen.uniform(50_000, seed=8)
    q = gen.uniform(7001, seed=9) * 1.4 - 0.2
    idx = KnnIndex(torch.from_numpy(x).to(_dev()))
    got = idx.knn_query(torch.from_numpy(q).to(_dev()), 12).cpu().numpy()
    assert bit_equal(got, bruteforce(x, queries=q, k=12))
    # the same index answers self queries afterwards
    got2 = idx.knn_self(5).cpu().numpy()
    sel = np.arange(0, 50_000, 50)
    assert bit_equal(got2[sel], bruteforce(x, queries=x[sel], k=5))
    idx.close()


def test_stats_counters():
    from mpi_cuda_largescaleknn_b200.ops.knn import KnnIndex
    x = gen.uniform(100_000, seed=10)
    idx = KnnIndex(torch.from_numpy(x).to(_dev()))
    idx.enable_stats(True)
    idx.knn_self(10)
    torch.cuda.synchronize()
    st = idx.stats()
    groups = (100_000 + 31) // 32
    assert st["leaves_visited"] >= groups            # at least its own leaf
    assert st["leaves_visited"] < groups * 400       # culling works
    assert st["heap_inserts"] >= 100_000 * 10
    idx.close()


def test_ring_world1_and_baseline_world1_equal_self():
    from mpi_cuda_largescaleknn_b200.ops.knn import KnnIndex
    from mpi_cuda_largescaleknn_b200.parallel.ring import FusedRing, NcclRingBaseline
    x = gen.uniform(40_000, seed=12)
    idx = KnnIndex(torch.from_numpy(x).to(_dev()))
    ref = idx.knn_self(20).cpu().numpy()
    ring = FusedRing(idx, 20, groups_per_chunk=100)      # 13 chunks
    assert bit_equal(ring.run().cpu().numpy(), ref)
    assert bit_equal(ring.run().cpu().numpy(), ref)      # second run: sequence numbers advance
    ring.close()
    base = NcclRingBaseline(idx, 20, groups_per_chunk=300)
    assert bit_equal(base.run().cpu().numpy(), ref)
    idx.close()


def test_against_plain_torch_fp32_brute_force():
    """An oracle that shares no code with the library (not even knn_math.h): squared distances from plain torch fp32
    elementwise arithmetic, k-th smallest by torch.topk. torch rounds x*x + y*y + z*z differently from the library's
    fmaf chain, so the comparison uses the tolerance of the reference's own verifyKNN
    (cudaKDTree/testing/floatN-knn-and-fcp.cu:315-323): relative 1e-6."""
    from mpi_cuda_largescaleknn_b200.ops.knn import KnnIndex
    for n, k, kind in ((6000, 10, "uniform"), (5000, 100, "uniform"), (7000, 32, "clustered")):
        x = gen.uniform(n, seed=n) if kind == "uniform" else gen.clustered_surface(n, seed=n)
        xd = torch.from_numpy(x).to(_dev())
        diff = xd[:, None, :] - xd[None, :, :]
        d2 = (diff * diff).sum(dim=2)
        want = torch.sqrt(torch.topk(d2, k, dim=1, largest=False).values[:, k - 1])
        idx = KnnIndex(xd)
        got = idx.knn_self(k)
        idx.close()
        rel = (got - want).abs() / torch.maximum(want, torch.tensor(1e-30, device=_dev()))
        assert float(rel.max().item()) <= 1e-6 or float((got - want).abs().max().item()) <= 1e-9, (n, k, float(rel.max().item()))


def test_planar_points_are_indexed_natively():
    """[N,2] points through b200knn_index_create_xy (no padding on the Python side): self kNN, external queries and
    find-closest-point equal the oracle run on the same points with z = 0."""
    from mpi_cuda_largescaleknn_b200.ops.knn import KnnIndex, fcp
    xy = gen.uniform(40_000, seed=41)[:, :2].copy()
    qy = gen.uniform(3_000, seed=42)[:, :2].copy()
    pad = lambda a: np.concatenate([a, np.zeros((len(a), 1), np.float32)], axis=1)
    idx = KnnIndex(torch.from_numpy(xy).to(_dev()))
    assert idx.dims == 2
    assert bit_equal(idx.knn_self(10).cpu().numpy(), bruteforce(pad(xy), k=10))
    assert bit_equal(idx.knn_query(torch.from_numpy(qy).to(_dev()), 7).cpu().numpy(), bruteforce(pad(xy), queries=pad(qy), k=7))
    ids, d = fcp(idx, torch.from_numpy(qy).to(_dev()))
    assert bit_equal(d.cpu().numpy(), bruteforce(pad(xy), queries=pad(qy), k=1))
    with pytest.raises(ValueError):
        idx.knn_query(torch.from_numpy(gen.uniform(10, seed=1)).to(_dev()), 3)      # 3-D queries against a planar index
    idx.close()


def test_fcp_search_params_eps_and_inspection_budget():
    """FcpSearchParams (cukd/fcp.h:35-52): eps > 0 may return a point up to sqrt(1 + eps) times farther than the
    nearest, never nearer; a finite inspection budget returns real points whose distance is an upper bound; with
    eps = 0 and no budget the result is the exact one."""
    from mpi_cuda_largescaleknn_b200.ops.knn import KnnIndex, fcp
    x = gen.clustered_surface(60_000, seed=31)
    q = gen.uniform(8_000, seed=32)
    xd, qd = torch.from_numpy(x).to(_dev()), torch.from_numpy(q).to(_dev())
    idx = KnnIndex(xd)
    want = bruteforce(x, queries=q, k=1)
    ids0, d0 = fcp(idx, qd, eps=0.0, far_node_inspect_budget=-1)
    assert bit_equal(d0.cpu().numpy(), want)
    for eps in (0.1, 1.0):
        ids, d = fcp(idx, qd, eps=eps)
        d, ids = d.cpu().numpy(), ids.cpu().numpy()
        assert np.all(d >= want) and np.all(d <= want * np.float32(math.sqrt(1.0 + eps)) * np.float32(1.0001))
        assert ids.min() >= 0
        real = np.linalg.norm(x[ids[:500]].astype(np.float64) - q[:500].astype(np.float64), axis=1)
        assert np.allclose(real, d[:500], rtol=1e-5, atol=1e-7)          # the reported point is at the reported distance
    for budget in (0, 4):
        ids, d = fcp(idx, qd, far_node_inspect_budget=budget)
        d, ids = d.cpu().numpy(), ids.cpu().numpy()
        found = ids >= 0
        assert np.all(d[found] >= want[found])
        real = np.linalg.norm(x[ids[found][:500]].astype(np.float64) - q[found][:500].astype(np.float64), axis=1)
        assert np.allclose(real, d[found][:500], rtol=1e-5, atol=1e-7)
    idx.close()


def test_fcp_returns_closest_point_id_and_distance():
    """find-closest-point (cukd/fcp.h): id of the nearest indexed point + its distance, cut-off radius."""
    from mpi_cuda_largescaleknn_b200.ops.knn import KnnIndex, fcp
    x = gen.uniform(30_000, seed=21)
    q = gen.uniform(5_000, seed=22) * 1.2 - 0.1
    idx = KnnIndex(torch.from_numpy(x).to(_dev()), id_offset=1000)
    ids, dist = fcp(idx, torch.from_numpy(q).to(_dev()))
    ids, dist = ids.cpu().numpy(), dist.cpu().numpy()
    want = bruteforce(x, queries=q, k=1)
    assert bit_equal(dist, want)
    assert ids.min() >= 1000 and ids.max() < 1000 + len(x)
    # the returned point really is at the reported distance (recomputed with the oracle on that single point)
    chk = np.array([bruteforce(x[i - 1000][None, :], queries=q[j][None, :], k=1)[0] for j, i in enumerate(ids[:300])], np.float32)
    assert bit_equal(chk, dist[:300])
    # cut-off: nothing within r -> id -1 and distance sqrt(r*r)
    far = np.full((4, 3), 50.0, np.float32)
    ids2, dist2 = fcp(idx, torch.from_numpy(far).to(_dev()), max_radius=0.5)
    assert np.all(ids2.cpu().numpy() == -1)
    assert np.all(dist2.cpu().numpy() == np.float32(0.5))
    idx.close()


def test_index_over_records_with_payload():
    """b200knn_index_create_records: positions read in place out of [N, F] records (payload columns before and after
    them). Self kNN, external queries and find-closest-point must give the bits of an index built on a contiguous
    copy of the positions, and the ids returned must be record numbers — the handle to the payload."""
    from mpi_cuda_largescaleknn_b200.ops.knn import KnnIndex, fcp
    for dims, first, width in ((3, 2, 7), (2, 1, 4), (3, 0, 3)):
        n = 30_000
        rec = torch.from_numpy(np.random.default_rng(dims * 10 + first).random((n, width), dtype=np.float32)).to(_dev())
        pos = rec[:, first:first + dims].contiguous()
        q = torch.from_numpy(np.random.default_rng(5).random((2_000, dims), dtype=np.float32)).to(_dev())
        a, b = KnnIndex.from_records(rec, first_column=first, dims=dims), KnnIndex(pos)
        assert torch.equal(a.knn_self(16), b.knn_self(16))
        assert torch.equal(a.knn_query(q, 8), b.knn_query(q, 8))
        ia, da = fcp(a, q)
        ib, db = fcp(b, q)
        assert torch.equal(ia, ib) and torch.equal(da, db)
        # the id is the record number: the record's own position is the closest point that was reported
        hit = rec[ia.long()][:, first:first + dims]
        d = torch.sqrt(((hit - q) ** 2).sum(dim=1))
        assert float((d - da).abs().max().item()) <= 1e-6
        a.close()
        b.close()


def test_index_build_falls_back_to_full_keys_on_heavy_cells():
    """More than 1024 points inside one cell of the key's leading 40 bits (here: 5000 points within 1e-6 of the
    extent, next to a uniform cloud): the build notices on the device and sorts again by all 64 bits. Results are
    exact either way; this checks that the switch happens and that ordinary clouds do not take it."""
    from mpi_cuda_largescaleknn_b200.ops.knn import KnnIndex
    x = gen.uniform(60_000, seed=5)
    tight = (np.float32(0.5) + np.random.default_rng(6).random((5000, 3), dtype=np.float32) * np.float32(1e-6)).astype(np.float32)
    both = np.concatenate([x, tight])
    a = KnnIndex(torch.from_numpy(x).to(_dev()))
    b = KnnIndex(torch.from_numpy(both).to(_dev()))
    assert a.sort_bits == 40 and b.sort_bits == 64
    sel = np.concatenate([np.arange(0, 60_000, 97), np.arange(60_000, 65_000, 41)])
    got = b.knn_self(20).cpu().numpy()
    assert bit_equal(got[sel], bruteforce(both, queries=both[sel], k=20))
    a.close()
    b.close()
