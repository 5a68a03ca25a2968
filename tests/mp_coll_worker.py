"""Multi-GPU worker: device-side all-gather / all-reduce over the symmetric heap against what
torch.distributed computes for the same values, many in a row (parity double-buffering, sequence
numbers), interleaved with kNN calls that grow the heap; then one culled-exchange case with the
device vote and with the host vote (B200KNN_HOST_VOTE)."""
import os
import sys
import time
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
    g = torch.Generator(device="cpu").manual_seed(1234)      # same stream of values on every rank

    def knn_case(N, k, mode):
        nonlocal fails
        x_all = gen.uniform(N, seed=N)
        b, e = global_slice(N, rank, world)
        x = x_all[b:e]
        sel = np.arange(0, len(x), max(1, len(x) // 800))
        ref = bruteforce(x_all, queries=x[sel], k=k)
        got = eng.knn(torch.from_numpy(x).to(dev), k, mode=mode, global_begin=b).cpu().numpy()
        if not bit_equal(got[sel], ref):
            fails += 1
            print(f"[rank {rank}] MISMATCH {mode} N={N}", flush=True)

    skew = np.random.default_rng(1000 + rank)                # different on every rank: launch skew
    for it in range(40):
        time.sleep(float(skew.random()) * 0.003)
        words = int(torch.randint(1, 33, (1,), generator=g))
        table = torch.randint(-2**31, 2**31 - 1, (world, words), generator=g, dtype=torch.int64).to(torch.int32)
        got = eng.allgather_dev(table[rank].to(dev))
        if not torch.equal(got.cpu(), table):
            fails += 1
            print(f"[rank {rank}] allgather_dev mismatch at it={it} words={words}", flush=True)
        count = int(torch.randint(1, 17, (1,), generator=g))
        t64 = torch.randint(-2**58, 2**58, (world, count), generator=g, dtype=torch.int64)
        for op, want in (("max", t64.max(0).values), ("sum", t64.sum(0)), ("min", t64.min(0).values)):
            got = eng.allreduce_dev(t64[rank].to(dev), op)
            if not torch.equal(got.cpu(), want):
                fails += 1
                print(f"[rank {rank}] allreduce_dev {op} mismatch at it={it} count={count}", flush=True)
        if it == 5:
            knn_case(40_000, 16, "prepartitioned")           # grows the heap: flags restart at zero
        if it == 20:
            knn_case(120_000, 100, "redistribute")
    # back-to-back without host synchronisation in between: only stream order separates them
    acc = torch.full((4,), rank + 1, dtype=torch.int64, device=dev)
    for i in range(2000):
        if i % 97 == rank:
            time.sleep(float(skew.random()) * 0.002)
        acc = eng.allreduce_dev(acc, "sum") % 1_000_003
    want = torch.full((4,), 0, dtype=torch.int64)
    vals = [r + 1 for r in range(world)]
    for _ in range(2000):
        s = sum(vals) % 1_000_003
        vals = [s] * world
    want[:] = vals[0]
    if not torch.equal(acc.cpu(), want):
        fails += 1
        print(f"[rank {rank}] chained allreduce_dev mismatch: {acc.cpu().tolist()} want {want.tolist()}", flush=True)

    os.environ["B200KNN_HOST_VOTE"] = "1"
    knn_case(70_000, 32, "prepartitioned")
    del os.environ["B200KNN_HOST_VOTE"]
    knn_case(70_000, 32, "prepartitioned")

    t = torch.tensor([fails], device=dev)
    dist.all_reduce(t)
    if rank == 0:
        print("MP_COLL_OK" if int(t.item()) == 0 else f"MP_COLL_FAILED {int(t.item())}", flush=True)
    eng.close()
    dist.barrier()
    dist.destroy_process_group()
    return 0 if int(t.item()) == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
