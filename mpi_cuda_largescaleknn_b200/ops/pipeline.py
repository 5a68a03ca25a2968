"""Host-to-device prefetch for loops that feed point clouds from pinned host memory.

A kNN step reads its whole input before it can start (the bounding box comes first), so a single step
cannot overlap its own upload. A LOOP can: while step i computes, step i+1's cloud crosses PCIe on a
second stream into the other of two staging buffers. At the headline size that hides 57 ms of copy behind
a 1 s step. The reference has nothing comparable (it reads its file once, synchronously:
/root/reference/unorderedDataVariant.cu:41-63,153-168).

    pipe = HostPrefetcher(n_points, device)
    for staged in pipe.feed(pinned_clouds):        # staged: [n,3] CUDA tensor, ready on the current stream
        d = knn(staged, k)                         # or DistKnn.knn(staged, ...)

The way back is the mirror image (`ResultDrain`): step i writes its distances into one of two device buffers,
and they cross PCIe into pinned host memory on the copy stream while step i+1 computes; the host gets step i's
array when it hands in step i+1. With eight ranks reading 134 MB each back at the same moment the download is
several milliseconds of a 105 ms step, so it is worth hiding.

    drain = ResultDrain(n_points, device)
    for staged in pipe.feed(pinned_clouds):
        knn(staged, k, out=drain.next_out())
        prev = drain.submit()                      # pinned host tensor with the PREVIOUS step's distances (or None)
    last = drain.finish()
"""
from __future__ import annotations

from typing import Iterable, Iterator

import torch


class HostPrefetcher:
    """Two device staging buffers and a copy stream; ``feed`` yields each cloud once it is on the device."""

    def __init__(self, n_points: int, device: torch.device, dtype: torch.dtype = torch.float32):
        self.device = torch.device(device)
        self.stage = [torch.empty((int(n_points), 3), dtype=dtype, device=self.device) for _ in range(2)]
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self.ready = [torch.cuda.Event(), torch.cuda.Event()]    # upload into slot finished
        self.free = [torch.cuda.Event(), torch.cuda.Event()]     # computation that used slot finished

    def _upload(self, slot: int, host: torch.Tensor) -> None:
        if not host.is_pinned():
            raise ValueError("HostPrefetcher needs pinned host tensors (torch.empty(..., pin_memory=True))")
        if tuple(host.shape) != tuple(self.stage[slot].shape):
            raise ValueError(f"expected a {tuple(self.stage[slot].shape)} cloud, got {tuple(host.shape)}")
        with torch.cuda.stream(self.copy_stream):
            # the computation that last read this slot must be done (no-op the first time round)
            self.copy_stream.wait_event(self.free[slot])
            self.stage[slot].copy_(host, non_blocking=True)
            self.ready[slot].record(self.copy_stream)

    def feed(self, host_clouds: Iterable[torch.Tensor]) -> Iterator[torch.Tensor]:
        """Yield device copies of ``host_clouds`` in order. The consumer must finish using a yielded tensor
        (on the current stream) before asking for the next one; the upload of the next cloud is already in
        flight by then."""
        main = torch.cuda.current_stream(self.device)
        it = iter(host_clouds)
        cur = next(it, None)
        if cur is None:
            return
        self._upload(0, cur)
        i = 0
        while cur is not None:
            nxt = next(it, None)
            if nxt is not None:
                self._upload((i + 1) % 2, nxt)
            slot = i % 2
            main.wait_event(self.ready[slot])
            yield self.stage[slot]
            self.free[slot].record(main)
            i += 1
            cur = nxt


class ResultDrain:
    """Two device result buffers, two pinned host buffers and a copy stream: the download of step i's result
    overlaps step i+1. ``next_out()`` -> device buffer for the step about to be enqueued; ``submit()`` after the
    step has been enqueued -> the previous step's result on the host, complete; ``finish()`` -> the last one."""

    def __init__(self, n_points: int, device: torch.device, dtype: torch.dtype = torch.float32):
        self.device = torch.device(device)
        self.out = [torch.empty(int(n_points), dtype=dtype, device=self.device) for _ in range(2)]
        self.host = [torch.empty(int(n_points), dtype=dtype, pin_memory=True) for _ in range(2)]
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self.filled = [torch.cuda.Event(), torch.cuda.Event()]   # the step that writes the slot has finished
        self.done = [torch.cuda.Event(), torch.cuda.Event()]     # the slot has arrived in host memory
        self.steps = 0

    def next_out(self) -> torch.Tensor:
        slot = self.steps % 2
        # the download that last read this slot must be done before the next step overwrites it
        torch.cuda.current_stream(self.device).wait_event(self.done[slot])
        return self.out[slot]

    def submit(self):
        slot = self.steps % 2
        self.filled[slot].record(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(self.filled[slot])
            self.host[slot].copy_(self.out[slot], non_blocking=True)
            self.done[slot].record(self.copy_stream)
        self.steps += 1
        if self.steps < 2:
            return None
        prev = self.steps % 2                 # = slot of step (steps - 2)
        self.done[prev].synchronize()
        return self.host[prev]

    def finish(self):
        if self.steps == 0:
            return None
        last = (self.steps - 1) % 2
        self.done[last].synchronize()
        return self.host[last]
