"""Single-GPU ops: spatial index + k-th-nearest-neighbour distance queries on torch CUDA tensors.

``knn(points, k)`` is the thin PyTorch entry point the north-star asks for: it takes an ``[N,3]``
float32 CUDA tensor and returns a float32 ``[N]`` tensor with the distance from every point to its
k-th nearest neighbour (the point itself included), exactly what the reference's
``runQuery`` + ``extractFinalResult`` pair produces (/root/reference/unorderedDataVariant.cu:75-103).
All work runs in the hand-written sm_100a kernels of libb200knn; torch only owns the tensors.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Optional

import numpy as np
import torch

from ._lib import check, lib


def _stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def _as_xyz(points: torch.Tensor, dims: int = 0) -> torch.Tensor:
    """[N,3] (or, for a planar index, [N,2]) float32 contiguous. With ``dims`` = 0 an [N,2] input is padded with
    z = 0 (entry points that only take float3); with ``dims`` = 2 or 3 the shape must match the index."""
    if points.dim() != 2 or points.shape[1] not in (2, 3):
        raise ValueError("points must have shape [N, 3] (or [N, 2])")
    if not points.is_cuda:
        raise ValueError("points must be a CUDA tensor (use utils.oracle for CPU work)")
    points = points.detach().to(torch.float32)
    if dims in (2, 3):
        if points.shape[1] != dims:
            raise ValueError(f"this index holds {dims}-D points; got [N, {points.shape[1]}]")
    elif points.shape[1] == 2:
        points = torch.cat([points, points.new_zeros((points.shape[0], 1))], dim=1)
    return points.contiguous()


class KnnIndex:
    """Morton-sorted point set + 32-ary box tree on the current CUDA device.

    Role of ``cukd::buildTree`` (cudaKDTree/cukd/builder.h:108-151). The index keeps its own sorted
    copy of the points (16 B/point) so the caller's tensor may be freed afterwards.
    """

    def __init__(self, points: torch.Tensor, ids: Optional[torch.Tensor] = None, id_offset: int = 0):
        self.dims = int(points.shape[1]) if points.dim() == 2 and points.shape[1] in (2, 3) else 3
        xyz = _as_xyz(points, self.dims)          # [N,2] stays [N,2]: the C ABI indexes planar points natively
        self.device = xyz.device
        self.n = int(xyz.shape[0])
        self.id_offset = int(id_offset)
        self._h = C.c_void_p()
        ids_ptr = None
        if ids is not None:
            ids = ids.to(device=self.device, dtype=torch.int32).contiguous()
            ids_ptr = ids.data_ptr()
        with torch.cuda.device(self.device):
            create = lib().b200knn_index_create
            if self.dims == 2:
                create = lib().b200knn_index_create_xy
                create.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p]
                create.restype = C.c_int
            check(create(xyz.data_ptr() if self.n else None, self.n, ids_ptr,
                         self.id_offset, _stream_ptr(), C.byref(self._h)),
                  "index_create")

    @classmethod
    def from_records(cls, records: torch.Tensor, first_column: int = 0, dims: int = 3,
                     ids: Optional[torch.Tensor] = None, id_offset: int = 0) -> "KnnIndex":
        """Index the positions embedded in caller records without copying them out first: ``records`` is a
        contiguous float32 CUDA tensor ``[N, F]`` (position + any payload columns), the position is ``dims`` (2 or
        3) columns starting at ``first_column``. Role of the reference's data traits
        (cudaKDTree/cukd/data.h:56-112): ids returned by ``fcp`` / ``neighbors`` are row numbers, so the payload
        of a hit is ``records[id]``. Queries are plain ``[nq, dims]`` tensors."""
        if records.dim() != 2 or records.dtype != torch.float32 or not records.is_contiguous():
            raise ValueError("from_records needs a contiguous float32 tensor [N, F]")
        if dims not in (2, 3) or first_column < 0 or first_column + dims > records.shape[1]:
            raise ValueError("the position columns do not fit inside a record")
        self = cls.__new__(cls)
        self.dims = int(dims)
        self.device = records.device
        self.n = int(records.shape[0])
        self.id_offset = int(id_offset)
        self._h = C.c_void_p()
        ids_ptr = None
        if ids is not None:
            ids = ids.to(device=self.device, dtype=torch.int32).contiguous()
            ids_ptr = ids.data_ptr()
        create = lib().b200knn_index_create_records
        create.argtypes = [C.c_void_p, C.c_int64, C.c_size_t, C.c_size_t, C.c_int, C.c_void_p, C.c_uint32, C.c_void_p,
                           C.c_void_p]
        create.restype = C.c_int
        with torch.cuda.device(self.device):
            check(create(records.data_ptr() if self.n else None, self.n, int(records.shape[1]) * 4, int(first_column) * 4,
                         self.dims, ids_ptr, self.id_offset, _stream_ptr(), C.byref(self._h)), "index_create_records")
        return self

    def close(self) -> None:
        if getattr(self, "_h", None) is not None and self._h.value:
            lib().b200knn_index_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def handle(self) -> C.c_void_p:
        return self._h

    @property
    def num_levels(self) -> int:
        return int(lib().b200knn_index_num_levels(self._h))

    @property
    def sort_bits(self) -> int:
        """Leading bits of the 63-bit Morton key the build sorted by (40; 64 after the heavy-cell fallback)."""
        fn = lib().b200knn_index_sort_bits
        fn.argtypes = [C.c_void_p]
        fn.restype = C.c_int
        return int(fn(self._h))

    def bounds(self) -> np.ndarray:
        out = np.zeros(6, np.float32)
        check(lib().b200knn_index_bounds(self._h, out.ctypes.data), "index_bounds")
        return out

    def sorted_points(self) -> np.ndarray:
        """Host copy of the sorted points as ``[n,4]`` float32 (w = id bits)."""
        out = np.zeros((self.n, 4), np.float32)
        check(lib().b200knn_index_copy_sorted(self._h, out.ctypes.data), "index_copy_sorted")
        return out

    def boxes(self, level: int) -> np.ndarray:
        cnt = C.c_int(0)
        check(lib().b200knn_index_copy_boxes(self._h, level, None, C.byref(cnt)), "index_copy_boxes")
        out = np.zeros((cnt.value, 8), np.float32)
        check(lib().b200knn_index_copy_boxes(self._h, level, out.ctypes.data, C.byref(cnt)), "index_copy_boxes")
        return out

    def enable_stats(self, on: bool = True) -> None:
        check(lib().b200knn_index_enable_stats(self._h, int(on)))

    def stats(self) -> dict:
        out = np.zeros(8, np.uint64)
        check(lib().b200knn_index_stats(self._h, out.ctypes.data))
        return {"leaves_visited": int(out[0]), "candidates": int(out[1]),
                "heap_inserts": int(out[2]), "inner_nodes": int(out[3]),
                "dfs_fallbacks": int(out[4]), "purges": int(out[5]), "list_flushes": int(out[6]),
                "slowest_group_kcycles": int(out[7]) >> 32, "slowest_group": int(out[7]) & 0xffffffff,
                "slowest_group_did": self._slowest()}

    def _slowest(self) -> dict:
        out = np.zeros(8, np.uint64)
        fn = lib().b200knn_index_stats_slowest
        fn.argtypes = [C.c_void_p, C.c_void_p]
        fn.restype = C.c_int
        check(fn(self._h, out.ctypes.data))
        return {"leaves": int(out[0]), "inner": int(out[1]), "flushes": int(out[2]), "fallbacks": int(out[3]),
                "purges": int(out[4]), "lane0_candidates": int(out[5])}

    def knn_self(self, k: int, max_radius: float = math.inf, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """k-th-neighbour distance of every indexed point, in the caller's original order."""
        if out is None:
            out = torch.empty(self.n, dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            check(lib().b200knn_knn_self(self._h, int(k), float(max_radius), out.data_ptr(), _stream_ptr()),
                  "knn_self")
        return out

    def knn_query(self, queries: torch.Tensor, k: int, max_radius: float = math.inf) -> torch.Tensor:
        """k-th-neighbour distance of arbitrary query points against the indexed points."""
        q = _as_xyz(queries, self.dims)
        out = torch.empty(q.shape[0], dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            check(lib().b200knn_knn_query(self._h, q.data_ptr() if q.shape[0] else None, int(q.shape[0]),
                                          int(k), float(max_radius), out.data_ptr(), _stream_ptr()),
                  "knn_query")
        return out


def neighbors(index: "KnnIndex", queries: Optional[torch.Tensor], k: int, max_radius: float = math.inf,
              sort: bool = True):
    """The k nearest indexed points of every query: ``(ids int32 [nq,k], distances float32 [nq,k])``.

    ``queries=None`` asks for the neighbours of the indexed points themselves (self included), rows in
    the caller's point order. With ``sort=True`` (default) rows are ordered nearest first; the library
    itself returns them unordered. Missing neighbours (fewer than k points, or none within
    ``max_radius``) are id -1 / distance +inf. Role of ``cukd::knn`` with a candidate list that is read
    back (cudaKDTree/cukd/knn.h:311-326, 617-631)."""
    if queries is None:
        nq, qptr = index.n, None
    else:
        q = _as_xyz(queries, index.dims)
        nq, qptr = int(q.shape[0]), (q.data_ptr() if q.shape[0] else None)
        if nq == 0:
            return (torch.empty((0, k), dtype=torch.int32, device=index.device),
                    torch.empty((0, k), dtype=torch.float32, device=index.device))
    ids = torch.empty((nq, int(k)), dtype=torch.int32, device=index.device)
    dist = torch.empty((nq, int(k)), dtype=torch.float32, device=index.device)
    fn = lib().b200knn_knn_neighbors
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
    fn.restype = C.c_int
    with torch.cuda.device(index.device):
        check(fn(index.handle, qptr, nq, int(k), float(max_radius), ids.data_ptr(), dist.data_ptr(), _stream_ptr()),
              "knn_neighbors")
    if sort and nq:
        dist, order = torch.sort(dist, dim=1, stable=True)
        ids = torch.gather(ids, 1, order)
    return ids, dist


def verify_sample(points: torch.Tensor, result: torch.Tensor, k: int, samples: int = 64, seed: int = 0,
                  group=None, max_radius: float = math.inf) -> dict:
    """Check ``samples`` (<= 64) randomly chosen entries of a self-kNN ``result`` against ALL points with one
    counting pass on the GPU: ``result[i]`` is right iff fewer than k points are strictly closer than it and at
    least k are no farther. A result equal to ``sqrt(max_radius^2)`` (the cut-off) is right iff fewer than k
    points lie strictly inside the cut-off; ``+inf`` iff fewer than k points exist at all.
    On several ranks (``group``, or the default process group when one is initialised) every rank passes its own
    ``points`` / ``result`` slice and ALL ranks must call; the sample is drawn on the first rank that holds any
    points and checked against the union of all ranks' points.
    Returns ``{"queries": n, "violations": v}``. Independent of the index, the lists and the traversal."""
    import torch.distributed as dist
    xyz = _as_xyz(points)
    dev = xyz.device
    n_local = int(xyz.shape[0])
    multi = group is not None or (dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1)
    rank, src = 0, 0
    if multi:
        rank = dist.get_rank(group)
        counts = torch.zeros(dist.get_world_size(group), dtype=torch.int64, device=dev)
        counts[rank] = n_local
        dist.all_reduce(counts, group=group)
        holders = torch.nonzero(counts > 0).flatten()
        if holders.numel() == 0:
            return {"queries": 0, "violations": 0}
        src = int(holders[0].item())
        n_src = int(counts[src].item())
    else:
        if n_local == 0:
            return {"queries": 0, "violations": 0}
        n_src = n_local
    nq = max(1, min(int(samples), 64, n_src))
    q = torch.zeros((nq, 3), dtype=torch.float32, device=dev)
    r = torch.zeros(nq, dtype=torch.float32, device=dev)
    if rank == src:
        g = torch.Generator(device="cpu").manual_seed(seed)
        pick = torch.randint(0, n_local, (nq,), generator=g).to(dev)
        q.copy_(xyz[pick])
        r.copy_(result.to(torch.float32)[pick])
    if multi:
        gsrc = dist.get_global_rank(group, src) if group is not None else src
        dist.broadcast(q, src=gsrc, group=group)
        dist.broadcast(r, src=gsrc, group=group)
    less = torch.zeros(nq, dtype=torch.int64, device=dev)
    less_equal = torch.zeros(nq, dtype=torch.int64, device=dev)
    fn = lib().b200knn_count_within
    fn.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    fn.restype = C.c_int
    with torch.cuda.device(dev):
        check(fn(xyz.data_ptr() if n_local else None, n_local, q.data_ptr(), nq, r.data_ptr(),
                 less.data_ptr(), less_equal.data_ptr(), _stream_ptr()), "count_within")
    if multi:
        dist.all_reduce(less, group=group)
        dist.all_reduce(less_equal, group=group)
    exact = ((less >= k) | (less_equal < k)) & torch.isfinite(r)
    if math.isfinite(max_radius):
        # a result AT the cut-off says "fewer than k points strictly inside it" and nothing about ties
        cut = torch.sqrt(torch.tensor(float(max_radius), dtype=torch.float32) ** 2).to(dev)
        at_cut = r == cut
        exact = torch.where(at_cut, less >= k, exact)
    # +inf results claim "fewer than k points exist": then every point is counted as strictly closer
    bad = exact | (torch.isinf(r) & (less >= k))
    return {"queries": nq, "violations": int(bad.sum().item())}


class _FcpParams(C.Structure):
    _fields_ = [("max_radius", C.c_float), ("eps", C.c_float), ("far_node_inspect_budget", C.c_int32)]


def fcp(index: "KnnIndex", queries: torch.Tensor, max_radius: float = math.inf, eps: float = 0.0,
        far_node_inspect_budget: int = -1):
    """Find-closest-point: (ids int32 [nq], distances float32 [nq]) of the nearest indexed point of
    every query; id -1 where nothing lies within ``max_radius``. ``eps`` > 0 and ``far_node_inspect_budget`` >= 0
    make the search approximate (the reported point is a real point, possibly not the nearest). Role of
    ``cukd::fcp`` with ``FcpSearchParams`` (cukd/fcp.h:35-52)."""
    q = _as_xyz(queries, index.dims)
    ids = torch.empty(q.shape[0], dtype=torch.int32, device=index.device)
    dist = torch.empty(q.shape[0], dtype=torch.float32, device=index.device)
    prm = _FcpParams(float(max_radius), float(eps), int(far_node_inspect_budget))
    fn = lib().b200knn_fcp_query_ex
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(_FcpParams), C.c_void_p, C.c_void_p, C.c_void_p]
    fn.restype = C.c_int
    with torch.cuda.device(index.device):
        check(fn(index.handle, q.data_ptr() if q.shape[0] else None, int(q.shape[0]), C.byref(prm), ids.data_ptr(),
                 dist.data_ptr(), _stream_ptr()), "fcp_query_ex")
    return ids, dist


def knn(points: torch.Tensor, k: int, max_radius: float = math.inf) -> torch.Tensor:
    """Distance from each point to its k-th nearest neighbour among ``points`` (self included)."""
    idx = KnnIndex(points)
    try:
        return idx.knn_self(k, max_radius)
    finally:
        idx.close()
