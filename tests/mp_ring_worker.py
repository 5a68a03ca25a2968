"""Multi-GPU worker (launched by torchrun, one rank per GPU): fused ring and NCCL-baseline ring must
both reproduce the brute-force oracle over the UNION of all ranks' points, bit for bit."""
import math
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))

from mpi_cuda_largescaleknn_b200.ops.knn import KnnIndex  # noqa: E402
from mpi_cuda_largescaleknn_b200.parallel.ring import FusedRing, NcclRingBaseline  # noqa: E402
from mpi_cuda_largescaleknn_b200.utils import generators as gen  # noqa: E402
from mpi_cuda_largescaleknn_b200.utils.io import slice_bounds  # noqa: E402
from mpi_cuda_largescaleknn_b200.utils.oracle import bit_equal, bruteforce  # noqa: E402


def main() -> int:
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    fails = 0
    cases = [  # N, k, radius, groups_per_chunk, data
        (20_000, 10, math.inf, 0, "uniform"),
        (100_003, 100, math.inf, 64, "uniform"),       # many chunks, uneven slices, partial groups
        (50_001, 32, 0.05, 100, "uniform"),
        (60_000, 16, math.inf, 7, "clustered"),        # tiny chunks: many ring steps in flight at once
        (world * 3 + 1, 5, math.inf, 0, "uniform"),    # fewer points than a group per rank
    ]
    for N, k, r, gc, kind in cases:
        x_all = gen.uniform(N, seed=N) if kind == "uniform" else gen.clustered_surface(N, seed=N)
        b, e = slice_bounds(N, rank, world)
        x = x_all[b:e]
        idx = KnnIndex(torch.from_numpy(x).to(dev))
        sel = np.arange(0, len(x), max(1, len(x) // 1500))
        ref = bruteforce(x_all, queries=x[sel], k=k, max_radius=r)
        for name, cls in (("fused", FusedRing), ("nccl", NcclRingBaseline)):
            ring = cls(idx, k, groups_per_chunk=gc)
            for rep in range(2):   # second run exercises sequence-number advance
                got = ring.run(max_radius=r).cpu().numpy()
                ok = bit_equal(got[sel], ref)
                if not ok:
                    fails += 1
                    bad = np.flatnonzero(got[sel].view(np.uint32) != ref.view(np.uint32))
                    print(f"[rank {rank}] MISMATCH {name} N={N} k={k} r={r} gc={gc} rep={rep}: "
                          f"{len(bad)}/{len(sel)} differ, e.g. got {got[sel][bad[:3]]} want {ref[bad[:3]]}", flush=True)
            if hasattr(ring, "close"):
                ring.close()
        idx.close()
        if rank == 0:
            print(f"case N={N} k={k} r={r} gc={gc} {kind}: done", flush=True)
    t = torch.tensor([fails], device=dev)
    dist.all_reduce(t)
    if rank == 0:
        print("MP_RING_OK" if int(t.item()) == 0 else f"MP_RING_FAILED {int(t.item())}", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    return 0 if int(t.item()) == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
