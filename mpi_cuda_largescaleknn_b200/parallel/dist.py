"""Python face of the C++ distributed driver (csrc/dist.cu): one process per GPU.

``DistKnn`` hands the library a *bootstrap* — two tiny blocking host collectives implemented with
``torch.distributed`` (gloo if the process group has it, else NCCL through a staging tensor). The
library uses it for set-up (CUDA IPC handles, boxes, counts) and round fences; every byte of bulk
data moves inside the library's kernels over NVLink-mapped peer memory.

Modes (see include/b200knn.h):
  ``ring``            all-pairs fused ring (/root/reference/unorderedDataVariant.cu:173-205)
  ``prepartitioned``  points stay where they are, bbox-culled per-group exchange
                      (/root/reference/prePartitionedDataVariant.cu:304-357)
  ``redistribute``    unordered input scattered to spatial owners first, results returned afterwards
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Optional, Tuple

import torch
import torch.distributed as dist

from ..ops._lib import check, lib
from ..ops.knn import _as_xyz, _stream_ptr

MODES = {"ring": 0, "prepartitioned": 1, "redistribute": 2, "auto": 3}

ALLGATHER_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64)
BARRIER_FN = C.CFUNCTYPE(C.c_int, C.c_void_p)


class Bootstrap(C.Structure):
    _fields_ = [("ctx", C.c_void_p), ("rank", C.c_int), ("world", C.c_int),
                ("allgather", ALLGATHER_FN), ("barrier", BARRIER_FN)]


class DistReport(C.Structure):
    _fields_ = [("ms_total", C.c_double), ("ms_redistribute", C.c_double), ("ms_build", C.c_double),
                ("ms_query", C.c_double), ("ms_return", C.c_double), ("pending_groups", C.c_int64),
                ("points_after_redistribute", C.c_int64), ("chunks", C.c_int32), ("rounds", C.c_int32),
                ("ms_local_phase", C.c_double), ("ms_host_collectives", C.c_double), ("ms_rounds", C.c_double),
                ("auto_chose_redistribute", C.c_int32), ("reserved", C.c_int32)]

    def as_dict(self) -> dict:
        return {name: getattr(self, name) for name, _ in self._fields_}


def _ready() -> bool:
    return dist.is_available() and dist.is_initialized()


class TorchBootstrap:
    """allgather/barrier of host bytes over a torch.distributed process group."""

    def __init__(self, group=None):
        self.group = group
        if _ready():
            self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
            backend = str(dist.get_backend(group))
            self.device = torch.device("cpu") if "gloo" in backend else torch.device("cuda", torch.cuda.current_device())
        else:
            self.rank, self.world, self.device = 0, 1, torch.device("cpu")
        self._ag = ALLGATHER_FN(self._allgather)
        self._bar = BARRIER_FN(self._barrier)
        self.struct = Bootstrap(None, self.rank, self.world, self._ag, self._bar)

    def _allgather(self, ctx, send, recv, nbytes) -> int:
        try:
            payload = C.string_at(send, nbytes)
            mine = torch.frombuffer(bytearray(payload), dtype=torch.uint8).to(self.device)
            out = torch.empty(self.world * nbytes, dtype=torch.uint8, device=self.device)
            dist.all_gather_into_tensor(out, mine, group=self.group)
            raw = out.cpu().numpy().tobytes()
            C.memmove(recv, raw, len(raw))
            return 0
        except Exception as e:  # surfaced by the library as "bootstrap allgather failed"
            print(f"[b200knn bootstrap] allgather failed: {e!r}", flush=True)
            return 1

    def _barrier(self, ctx) -> int:
        try:
            if self.device.type == "cpu":
                dist.barrier(group=self.group)
            else:
                t = torch.zeros(1, device=self.device)
                dist.all_reduce(t, group=self.group)
                t.item()
            return 0
        except Exception as e:
            print(f"[b200knn bootstrap] barrier failed: {e!r}", flush=True)
            return 1


class DistKnn:
    """Reusable distributed engine (symmetric IPC heap is kept between calls)."""

    def __init__(self, group=None):
        self.bootstrap = TorchBootstrap(group)
        self.rank, self.world = self.bootstrap.rank, self.bootstrap.world
        self._h = C.c_void_p()
        fn = lib().b200knn_dist_create
        fn.argtypes = [C.POINTER(Bootstrap), C.POINTER(C.c_void_p)]
        fn.restype = C.c_int
        check(fn(C.byref(self.bootstrap.struct), C.byref(self._h)), "dist_create")
        self._knn = lib().b200knn_dist_knn
        self._knn.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int, C.c_float, C.c_int, C.c_int,
                              C.c_void_p, C.c_void_p, C.POINTER(DistReport)]
        self._knn.restype = C.c_int
        self.last_report: Optional[dict] = None

    @property
    def heap_bytes(self) -> int:
        fn = lib().b200knn_dist_heap_bytes
        fn.argtypes = [C.c_void_p]
        fn.restype = C.c_int64
        return int(fn(self._h))

    def knn(self, points: torch.Tensor, k: int, mode: str = "redistribute", global_begin: int = 0,
            max_radius: float = math.inf, groups_per_chunk: int = 0,
            out: Optional[torch.Tensor] = None) -> torch.Tensor:
        xyz = _as_xyz(points)
        n = int(xyz.shape[0])
        if out is None:
            out = torch.empty(n, dtype=torch.float32, device=xyz.device)
        rep = DistReport()
        with torch.cuda.device(xyz.device):
            check(self._knn(self._h, xyz.data_ptr() if n else None, n, int(global_begin), int(k), float(max_radius),
                            MODES[mode], int(groups_per_chunk), out.data_ptr() if n else None, _stream_ptr(),
                            C.byref(rep)), f"dist_knn[{mode}]")
        self.last_report = rep.as_dict()
        return out

    def allgather_dev(self, send: torch.Tensor) -> torch.Tensor:
        """Device-side all-gather of a small CUDA tensor (<= 128 bytes, 4-byte multiple): result is
        ``[world, *send.shape]``, produced by one kernel on the current stream (no host sync)."""
        send = send.contiguous()
        nbytes = send.numel() * send.element_size()
        out = torch.empty((self.world,) + tuple(send.shape), dtype=send.dtype, device=send.device)
        fn = lib().b200knn_dist_allgather_dev
        fn.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        fn.restype = C.c_int
        with torch.cuda.device(send.device):
            check(fn(self._h, send.data_ptr(), nbytes, out.data_ptr(), _stream_ptr()), "dist_allgather_dev")
        return out

    def allreduce_dev(self, send: torch.Tensor, op: str = "max") -> torch.Tensor:
        """Device-side element-wise all-reduce (max / sum / min) of up to 16 int64 values."""
        if send.dtype != torch.int64:
            raise TypeError("allreduce_dev works on int64 tensors")
        send = send.contiguous()
        out = torch.empty_like(send)
        fn = lib().b200knn_dist_allreduce_dev
        fn.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        fn.restype = C.c_int
        with torch.cuda.device(send.device):
            check(fn(self._h, send.data_ptr(), send.numel(), {"max": 0, "sum": 1, "min": 2}[op], out.data_ptr(),
                     _stream_ptr()), "dist_allreduce_dev")
        return out

    def close(self) -> None:
        if self._h.value:
            fn = lib().b200knn_dist_destroy
            fn.argtypes = [C.c_void_p]
            fn.restype = None
            fn(self._h)
            self._h = C.c_void_p()


def global_slice(total: int, rank: int, world: int) -> Tuple[int, int]:
    """The reference's even index split (/root/reference/unorderedDataVariant.cu:56-59)."""
    return total * rank // world, total * (rank + 1) // world
