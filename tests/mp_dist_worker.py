"""Multi-GPU worker: the C++ distributed driver in all three modes must reproduce the brute-force
oracle over the union of all ranks' points, bit for bit, and culling must actually cull."""
import math
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))

from mpi_cuda_largescaleknn_b200.parallel.dist import DistKnn, global_slice  # noqa: E402
from mpi_cuda_largescaleknn_b200.utils import generators as gen  # noqa: E402
from mpi_cuda_largescaleknn_b200.utils.oracle import bit_equal, bruteforce  # noqa: E402


def main() -> int:
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("cpu:gloo,cuda:nccl", device_id=dev)
    eng = DistKnn()
    fails = 0

    def check(name, got, ref, sel):
        nonlocal fails
        if not bit_equal(got[sel], ref):
            fails += 1
            bad = np.flatnonzero(got[sel].view(np.uint32) != ref.view(np.uint32))
            print(f"[rank {rank}] MISMATCH {name}: {len(bad)}/{len(sel)} differ, got {got[sel][bad[:3]]} want {ref[bad[:3]]}",
                  flush=True)

    # ---- unordered input: every mode, several shapes
    cases = [(20_000, 10, math.inf, 0, "uniform"), (100_003, 100, math.inf, 64, "uniform"),
             (50_001, 32, 0.05, 100, "uniform"), (60_000, 16, math.inf, 7, "clustered"),
             (world * 3 + 1, 5, math.inf, 0, "uniform"), (300_000, 100, math.inf, 0, "uniform"),
             (world * 1300, 1900, math.inf, 0, "uniform")]     # k beyond the shared-memory heap: global-heap kernels
    for N, k, r, gc, kind in cases:
        x_all = gen.uniform(N, seed=N) if kind == "uniform" else gen.clustered_surface(N, seed=N)
        b, e = global_slice(N, rank, world)
        x = x_all[b:e]
        xd = torch.from_numpy(x).to(dev)
        sel = np.arange(0, len(x), max(1, len(x) // 1200)) if len(x) else np.zeros(0, np.int64)
        ref = bruteforce(x_all, queries=x[sel], k=k, max_radius=r) if len(x) else np.zeros(0, np.float32)
        for mode in ("ring", "prepartitioned", "redistribute"):
            for rep in range(2):
                got = eng.knn(xd, k, mode=mode, global_begin=b, max_radius=r, groups_per_chunk=gc).cpu().numpy()
                check(f"{mode} N={N} k={k} r={r} gc={gc} rep={rep}", got, ref, sel)
        if rank == 0:
            print(f"case N={N} k={k} r={r} gc={gc} {kind}: done; last report {eng.last_report}", flush=True)

    # ---- pre-partitioned blobs: culling must leave most groups at home
    for overlap in (0.0, 0.10, 1.0):
        parts = gen.gaussian_blobs(40_000, world, overlap=overlap, seed=7)
        x_all = np.concatenate(parts)
        x = parts[rank]
        xd = torch.from_numpy(x).to(dev)
        sel = np.arange(0, len(x), 40)
        ref = bruteforce(x_all, queries=x[sel], k=20)
        got = eng.knn(xd, 20, mode="prepartitioned").cpu().numpy()
        check(f"blobs overlap={overlap}", got, ref, sel)
        pend_inplace = eng.last_report["pending_groups"]
        got = eng.knn(xd, 20, mode="auto").cpu().numpy()
        check(f"blobs overlap={overlap} auto", got, ref, sel)
        if overlap == 0.0 and eng.last_report["auto_chose_redistribute"] != 0:
            fails += 1
            print("AUTO FAILED: disjoint blobs must stay in place", flush=True)
        if overlap == 1.0 and eng.last_report["auto_chose_redistribute"] != 1:
            fails += 1
            print("AUTO FAILED: interpenetrating blobs must be redistributed", flush=True)
        eng.last_report["pending_groups"] = pend_inplace
        pend = torch.tensor([eng.last_report["pending_groups"]], device=dev)
        dist.all_reduce(pend)
        groups = world * ((40_000 + 31) // 32)
        if rank == 0:
            print(f"blobs overlap={overlap}: pending groups {int(pend.item())} of {groups}", flush=True)
        if overlap == 0.0 and int(pend.item()) > groups // 10:
            fails += 1
            print("CULLING FAILED: disjoint blobs should leave almost every group at home", flush=True)

    t = torch.tensor([fails], device=dev)
    dist.all_reduce(t)
    if rank == 0:
        print("MP_DIST_OK" if int(t.item()) == 0 else f"MP_DIST_FAILED {int(t.item())}", flush=True)
    eng.close()
    dist.barrier()
    dist.destroy_process_group()
    return 0 if int(t.item()) == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
