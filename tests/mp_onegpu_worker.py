"""Several ranks on ONE GPU (the reference's `-g` oversubscription, unorderedDataVariant.cu:122-123,138-143):
every rank uses device 0, torch.distributed runs on gloo only, the ranks' kernels talk through CUDA-IPC
mappings of each other's memory on the same device. The GPU time-slices between the processes, so the
in-kernel flag waits make progress a time slice at a time — slow, but it is the full multi-rank code
path (fused ring, culled exchange, redistribution, device collectives) on a single-GPU box.
Launched by tests/test_gpu_oversubscribed.py through torchrun."""
import math
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))

from mpi_cuda_largescaleknn_b200.ops.knn import KnnIndex  # noqa: E402
from mpi_cuda_largescaleknn_b200.parallel.dist import DistKnn, global_slice  # noqa: E402
from mpi_cuda_largescaleknn_b200.parallel.ring import FusedRing  # noqa: E402
from mpi_cuda_largescaleknn_b200.utils import generators as gen  # noqa: E402
from mpi_cuda_largescaleknn_b200.utils.oracle import bit_equal, bruteforce  # noqa: E402


def main() -> int:
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    dist.init_process_group("gloo")
    fails = 0

    def check(name, got, ref, sel):
        nonlocal fails
        if not bit_equal(got[sel], ref):
            fails += 1
            bad = np.flatnonzero(got[sel].view(np.uint32) != ref.view(np.uint32))
            print(f"[rank {rank}] MISMATCH {name}: {len(bad)}/{len(sel)} differ", flush=True)

    cases = [(20_000, 10, math.inf, 0, "uniform"), (30_003, 100, math.inf, 64, "uniform"),
             (25_001, 32, 0.05, 100, "uniform"), (24_000, 16, math.inf, 7, "clustered"),
             (world * 3 + 1, 5, math.inf, 0, "uniform")]

    # ---- fused ring (query.cu): in-kernel hand-off between processes on the same device
    for N, k, r, gc, kind in cases:
        x_all = gen.uniform(N, seed=N) if kind == "uniform" else gen.clustered_surface(N, seed=N)
        b, e = global_slice(N, rank, world)
        x = x_all[b:e]
        sel = np.arange(0, len(x), max(1, len(x) // 600))
        ref = bruteforce(x_all, queries=x[sel], k=k, max_radius=r)
        idx = KnnIndex(torch.from_numpy(x).to(dev))
        ring = FusedRing(idx, k, groups_per_chunk=gc)
        for rep in range(2):
            check(f"fused ring N={N} k={k} rep={rep}", ring.run(max_radius=r).cpu().numpy(), ref, sel)
        ring.close()
        idx.close()
    if rank == 0:
        print("ring cases done", flush=True)

    # ---- distributed driver (dist.cu), every mode, and the device collectives in between
    eng = DistKnn()
    g = torch.Generator(device="cpu").manual_seed(99)
    for ci, (N, k, r, gc, kind) in enumerate(cases):
        x_all = gen.uniform(N, seed=N) if kind == "uniform" else gen.clustered_surface(N, seed=N)
        b, e = global_slice(N, rank, world)
        x = x_all[b:e]
        xd = torch.from_numpy(x).to(dev)
        sel = np.arange(0, len(x), max(1, len(x) // 600))
        ref = bruteforce(x_all, queries=x[sel], k=k, max_radius=r)
        for mode in ("ring", "prepartitioned", "redistribute", "auto"):
            got = eng.knn(xd, k, mode=mode, global_begin=b, max_radius=r, groups_per_chunk=gc).cpu().numpy()
            check(f"{mode} N={N} k={k}", got, ref, sel)
        words = int(torch.randint(1, 33, (1,), generator=g))
        table = torch.randint(-2**31, 2**31 - 1, (world, words), generator=g, dtype=torch.int64).to(torch.int32)
        if not torch.equal(eng.allgather_dev(table[rank].to(dev)).cpu(), table):
            fails += 1
            print(f"[rank {rank}] allgather_dev mismatch (case {ci})", flush=True)
        t64 = torch.randint(-2**58, 2**58, (world, 5), generator=g, dtype=torch.int64)
        for op, want in (("max", t64.max(0).values), ("sum", t64.sum(0)), ("min", t64.min(0).values)):
            if not torch.equal(eng.allreduce_dev(t64[rank].to(dev), op).cpu(), want):
                fails += 1
                print(f"[rank {rank}] allreduce_dev {op} mismatch (case {ci})", flush=True)
        if rank == 0:
            print(f"dist case {ci} done: {eng.last_report}", flush=True)

    # ---- pre-partitioned blobs: disjoint clouds must mostly stay at home
    parts = gen.gaussian_blobs(12_000, world, overlap=0.0, seed=7)
    x_all = np.concatenate(parts)
    x = parts[rank]
    sel = np.arange(0, len(x), 40)
    got = eng.knn(torch.from_numpy(x).to(dev), 20, mode="prepartitioned").cpu().numpy()
    check("blobs", got, bruteforce(x_all, queries=x[sel], k=20), sel)
    pend = torch.tensor([eng.last_report["pending_groups"]])
    dist.all_reduce(pend)
    if int(pend.item()) > world * ((12_000 + 31) // 32) // 10:
        fails += 1
        print("CULLING FAILED: disjoint blobs should leave almost every group at home", flush=True)

    t = torch.tensor([fails])
    dist.all_reduce(t)
    if rank == 0:
        print("MP_ONEGPU_OK" if int(t.item()) == 0 else f"MP_ONEGPU_FAILED {int(t.item())}", flush=True)
    eng.close()
    dist.barrier()
    dist.destroy_process_group()
    return 0 if int(t.item()) == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
