"""File-level pipeline of the unordered variant in Python (same contract as the executable and as
/root/reference/unorderedDataVariant.cu:105-238): read this rank's slice of one flat float3 file,
run the distributed kNN, write this rank's slice of the single float32 output file.

    torchrun --nnodes=1 --nproc-per-node=8 --master-addr 127.0.0.1 \
        -m mpi_cuda_largescaleknn_b200.models.unordered points.float3 -o dist.float -k 100 [-r R] [--mode ring]
"""
from __future__ import annotations

import argparse
import math
import os
import sys

import torch
import torch.distributed as dist

from ..parallel.dist import DistKnn
from ..utils import io as kio


def run_unordered(in_path: str, out_path: str, k: int, max_radius: float = math.inf, mode: str = "redistribute",
                  engine: DistKnn | None = None) -> dict:
    rank = dist.get_rank() if dist.is_initialized() else 0
    world = dist.get_world_size() if dist.is_initialized() else 1
    pts, begin, total = kio.read_float3(in_path, rank, world)
    print(f"#{rank}/{world}: got {len(pts)} points to work on", flush=True)
    own = engine is None
    eng = engine or DistKnn()
    dev = torch.device("cuda", torch.cuda.current_device())
    x = torch.from_numpy(pts).to(dev)
    res = eng.knn(x, k, mode=mode, global_begin=begin, max_radius=max_radius).cpu().numpy()
    if rank == 0:
        print("done all queries...", flush=True)
        kio.write_floats_at(out_path, res, begin, total, create=True)
    if world > 1:
        dist.barrier()
    if rank != 0:
        kio.write_floats_at(out_path, res, begin, total, create=False)
    if world > 1:
        dist.barrier()
    report = dict(eng.last_report or {})
    if own:
        eng.close()
    return report


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("input")
    ap.add_argument("-o", dest="output", required=True)
    ap.add_argument("-k", type=int, required=True)
    ap.add_argument("-r", dest="max_radius", type=float, default=math.inf)
    ap.add_argument("--mode", default="redistribute", choices=["redistribute", "ring", "prepartitioned"])
    a = ap.parse_args(argv)
    if a.k < 1:
        ap.error("no k specified, or invalid k value")
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("cpu:gloo,cuda:nccl", device_id=torch.device("cuda", local))
    run_unordered(a.input, a.output, a.k, a.max_radius, a.mode)
    if dist.is_initialized():
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
