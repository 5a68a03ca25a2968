"""File-level pipeline of the pre-partitioned variant in Python (same contract as the executable and
as /root/reference/prePartitionedDataVariant.cu:176-388): one input file per rank (listed in a text
file), one output file per rank named ``<prefix>_%06d.float``.

    torchrun --nnodes=1 --nproc-per-node=8 --master-addr 127.0.0.1 \
        -m mpi_cuda_largescaleknn_b200.models.prepartitioned files.txt -o prefix -k 100 [-r R] [--mode auto]
"""
from __future__ import annotations

import argparse
import math
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

from ..parallel.dist import DistKnn
from ..utils import io as kio


def run_prepartitioned(list_path: str, out_prefix: str, k: int, max_radius: float = math.inf, mode: str = "auto",
                       engine: DistKnn | None = None) -> dict:
    rank = dist.get_rank() if dist.is_initialized() else 0
    world = dist.get_world_size() if dist.is_initialized() else 1
    names = kio.read_file_list(list_path)
    if len(names) != world:
        raise RuntimeError("number of input files does not match number of ranks")
    pts, _, _ = kio.read_float3(names[rank], 0, 1)
    lo = pts.min(0) if len(pts) else np.full(3, np.inf)
    hi = pts.max(0) if len(pts) else np.full(3, -np.inf)
    print(f"#{rank}/{world}: got {len(pts)} points to work on, bounds is ({lo[0]},{lo[1]},{lo[2]})-({hi[0]},{hi[1]},{hi[2]})",
          flush=True)
    own = engine is None
    eng = engine or DistKnn()
    dev = torch.device("cuda", torch.cuda.current_device())
    res = eng.knn(torch.from_numpy(pts).to(dev), k, mode=mode, max_radius=max_radius).cpu().numpy()
    if rank == 0:
        print("done all queries...", flush=True)
    np.ascontiguousarray(res, dtype="<f4").tofile(kio.per_rank_output_name(out_prefix, rank))
    if world > 1:
        dist.barrier()
    report = dict(eng.last_report or {})
    if own:
        eng.close()
    return report


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("file_list")
    ap.add_argument("-o", dest="prefix", required=True)
    ap.add_argument("-k", type=int, required=True)
    ap.add_argument("-r", dest="max_radius", type=float, default=math.inf)
    ap.add_argument("--mode", default="auto", choices=["auto", "prepartitioned", "ring", "redistribute"])
    a = ap.parse_args(argv)
    if a.k < 1:
        ap.error("no k specified, or invalid k value")
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("cpu:gloo,cuda:nccl", device_id=torch.device("cuda", local))
    run_prepartitioned(a.file_list, a.prefix, a.k, a.max_radius, "auto" if a.mode == "redistribute" else a.mode)
    if dist.is_initialized():
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
