"""Multi-GPU ring for unordered data: query groups + candidate heaps circulate, trees stay home.

Two transports over the same persistent kernel (csrc/query.cu):

``FusedRing``      the product. One kernel launch per kNN run per GPU; the kernel itself stores each
                   group's heap into the next GPU's inbox over NVLink (CUDA-IPC-mapped peer memory)
                   and hand-shakes with system-scope release/acquire flags. torch.distributed is
                   used only as plumbing: all-gather of 256-byte IPC blobs at set-up and one barrier
                   before/after a run.
``NcclRingBaseline`` the baseline the north-star says to beat: one kernel launch per ring step and
                   NCCL send/recv of the state buffers in between (what a CUDA-aware MPI ring would
                   do at best). Same results, no overlap.

Role of the MPI ring in /root/reference/unorderedDataVariant.cu:173-205, with the dataflow
reversed (there the tree travels and n*k*8 bytes of heaps stay resident for all R rounds).
"""
from __future__ import annotations

import ctypes as C
import math
from typing import List, Optional

import torch
import torch.distributed as dist

from ..ops._lib import IPC_BLOB_BYTES, StepArgs, check, lib
from ..ops.knn import KnnIndex, _stream_ptr


def _dist_ready() -> bool:
    return dist.is_available() and dist.is_initialized()


def rank_world(group=None):
    if not _dist_ready():
        return 0, 1
    return dist.get_rank(group), dist.get_world_size(group)


def _comm_device(group=None) -> torch.device:
    backend = dist.get_backend(group)
    return torch.device("cuda", torch.cuda.current_device()) if "nccl" in str(backend) else torch.device("cpu")


def allgather_bytes(payload: bytes, group=None) -> List[bytes]:
    """All-gather equal-sized byte strings (IPC blobs, counts) over the process group."""
    if not _dist_ready():
        return [payload]
    dev = _comm_device(group)
    world = dist.get_world_size(group)
    mine = torch.frombuffer(bytearray(payload), dtype=torch.uint8).to(dev)
    out = torch.empty(world * len(payload), dtype=torch.uint8, device=dev)
    dist.all_gather_into_tensor(out, mine, group=group)
    raw = out.cpu().numpy().tobytes()
    return [raw[i * len(payload):(i + 1) * len(payload)] for i in range(world)]


def allreduce_max_int(v: int, group=None) -> int:
    if not _dist_ready():
        return v
    t = torch.tensor([v], dtype=torch.int64, device=_comm_device(group))
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return int(t.item())


def barrier(group=None) -> None:
    if _dist_ready():
        dist.barrier(group=group)


class FusedRing:
    """In-kernel ring pass over NVLink peer memory (see module docstring)."""

    def __init__(self, index: KnnIndex, k: int, group=None, groups_per_chunk: int = 0):
        self.index = index
        self.k = int(k)
        self.group = group
        self.rank, self.world = rank_world(group)
        self.max_n = allreduce_max_int(index.n, group)
        self._h = C.c_void_p()
        check(lib().b200knn_ring_create(index.handle, self.rank, self.world, self.k, self.max_n,
                                        int(groups_per_chunk), C.byref(self._h)), "ring_create")
        blob = (C.c_ubyte * IPC_BLOB_BYTES)()
        check(lib().b200knn_ring_export(self._h, blob), "ring_export")
        blobs = b"".join(allgather_bytes(bytes(blob), group))
        check(lib().b200knn_ring_connect(self._h, blobs), "ring_connect")
        barrier(group)

    @property
    def comm_bytes(self) -> int:
        return int(lib().b200knn_ring_comm_bytes(self._h))

    def launch(self, max_radius: float = math.inf) -> None:
        """Enqueue the fused kernel on the current stream (no sync)."""
        check(lib().b200knn_ring_run(self._h, float(max_radius), _stream_ptr()), "ring_run")

    def collect(self, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Wait for every rank's kernel, then put this rank's results into original order."""
        torch.cuda.current_stream().synchronize()
        barrier(self.group)
        if out is None:
            out = torch.empty(self.index.n, dtype=torch.float32, device=self.index.device)
        check(lib().b200knn_ring_results(self._h, out.data_ptr(), _stream_ptr()), "ring_results")
        return out

    def run(self, max_radius: float = math.inf, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        self.launch(max_radius)
        return self.collect(out)

    def close(self) -> None:
        if self._h.value:
            torch.cuda.synchronize()
            barrier(self.group)
            check(lib().b200knn_ring_disconnect(self._h), "ring_disconnect")
            barrier(self.group)          # every peer has unmapped before anybody frees
            lib().b200knn_ring_destroy(self._h)
            self._h = C.c_void_p()


class NcclRingBaseline:
    """Step-per-launch ring with NCCL send/recv between steps (baseline transport)."""

    def __init__(self, index: KnnIndex, k: int, group=None, groups_per_chunk: int = 0):
        self.index, self.k, self.group = index, int(k), group
        self.rank, self.world = rank_world(group)
        self.max_n = allreduce_max_int(index.n, group)
        max_groups = max(1, (self.max_n + 31) // 32)
        gc = groups_per_chunk if groups_per_chunk > 0 else 65536
        self.gc = min(gc, max_groups)
        self.chunks = (max_groups + self.gc - 1) // self.gc
        dev = index.device
        mk = lambda *s, dt=torch.float32: torch.zeros(*s, dtype=dt, device=dev)
        self.in_heap = [mk(self.gc, self.k, 32) for _ in range(2)]
        self.in_pos = [mk(self.gc, 32, 4) for _ in range(2)]
        self.in_cnt = [mk(self.gc, dt=torch.int32) for _ in range(2)]
        self.out_heap = [mk(self.gc, self.k, 32) for _ in range(2)]
        self.out_pos = [mk(self.gc, 32, 4) for _ in range(2)]
        self.out_cnt = [mk(self.gc, dt=torch.int32) for _ in range(2)]
        self.res_stage = mk(max_groups * 32)     # results of groups finishing here (belong to `next`)
        self.res_home = mk(max_groups * 32)      # my own groups' results (sent by `prev`)

    def _args(self, step: int, max_radius: float) -> StepArgs:
        a = StepArgs()
        a.k, a.world, a.rank = self.k, self.world, self.rank
        a.groups_per_chunk, a.num_chunks = self.gc, self.chunks
        a.step_begin, a.step_end = step, step + 1
        a.max_radius = max_radius
        for b in range(2):
            a.inbox_heap[b] = self.in_heap[b].data_ptr()
            a.inbox_pos[b] = self.in_pos[b].data_ptr()
            a.inbox_cnt[b] = self.in_cnt[b].data_ptr()
            a.outbox_heap[b] = self.out_heap[b].data_ptr()
            a.outbox_pos[b] = self.out_pos[b].data_ptr()
            a.outbox_cnt[b] = self.out_cnt[b].data_ptr()
        a.result_sorted = (self.res_stage if self.world > 1 else self.res_home).data_ptr()
        return a

    def _exchange(self, send: List[torch.Tensor], recv: List[torch.Tensor]) -> None:
        nxt, prv = (self.rank + 1) % self.world, (self.rank - 1) % self.world
        ops = []
        for s, r in zip(send, recv):
            ops.append(dist.P2POp(dist.isend, s, nxt, group=self.group))
            ops.append(dist.P2POp(dist.irecv, r, prv, group=self.group))
        for w in dist.batch_isend_irecv(ops):
            w.wait()

    def run(self, max_radius: float = math.inf, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        R = self.world
        for step in range(self.chunks * R):
            hop, chunk, par = step % R, step // R, step & 1
            a = self._args(step, float(max_radius))
            check(lib().b200knn_run_steps(self.index.handle, C.byref(a), _stream_ptr()), "run_steps")
            if R == 1:
                continue
            op = par ^ 1
            if hop < R - 1:
                self._exchange([self.out_heap[op], self.out_pos[op], self.out_cnt[op]],
                               [self.in_heap[op], self.in_pos[op], self.in_cnt[op]])
            else:
                lo, hi = chunk * self.gc * 32, min((chunk + 1) * self.gc * 32, self.res_stage.numel())
                self._exchange([self.res_stage[lo:hi]], [self.res_home[lo:hi]])
        if out is None:
            out = torch.empty(self.index.n, dtype=torch.float32, device=self.index.device)
        check(lib().b200knn_index_unsort(self.index.handle, self.res_home.data_ptr(), out.data_ptr(),
                                         _stream_ptr()), "index_unsort")
        return out
