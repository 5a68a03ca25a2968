#!/usr/bin/env python
"""Why is one rank's local phase slower on uniform data? Rebuild each rank's share of the 256 Mi bench
input on ONE GPU (same generators as bench.py), and time/count its local all-kNN pass."""
import json, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from mpi_cuda_largescaleknn_b200.ops.knn import KnnIndex
from mpi_cuda_largescaleknn_b200.utils.generators import torch_uniform

dev = torch.device("cuda", 0)
N, world = 256 * 1024 * 1024, 8
parts = []
for r in range(world):
    b, e = N * r // world, N * (r + 1) // world
    parts.append(torch_uniform(e - b, dev, seed=1234 + r))
x = torch.cat(parts)
del parts
lo, hi = x.min(0).values, x.max(0).values
mid = (lo + hi) / 2
octant = ((x[:, 0] >= mid[0]).long() << 2) | ((x[:, 1] >= mid[1]).long() << 1) | (x[:, 2] >= mid[2]).long()
for r in (int(a) for a in (sys.argv[1:] or ["0", "2", "4", "7"])):
    pts = x[octant == r].contiguous()
    out = torch.empty(pts.shape[0], device=dev)
    idx = KnnIndex(pts)
    idx.knn_self(100, out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); idx.knn_self(100, out=out); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    idx.enable_stats(True); idx.knn_self(100, out=out); torch.cuda.synchronize()
    st = idx.stats(); g = (pts.shape[0] + 31) // 32
    print(json.dumps({"octant": r, "n": pts.shape[0], "levels": idx.num_levels, "ms": ms,
                      "leaves_per_group": st["leaves_visited"] / g, "inserts_per_query": st["heap_inserts"] / pts.shape[0],
                      "inner_per_group": st["inner_nodes"] / g, "fallbacks_per_group": st["dfs_fallbacks"] / g}), flush=True)
    idx.close(); del pts, out

# ---- follow-up: is it the VALUES? shift octant 2 into octant 0's coordinate range and vice versa
import pynvml
pynvml.nvmlInit(); h = pynvml.nvmlDeviceGetHandleByIndex(0)
for r, shift in ((2, -0.5), (0, +0.5), (2, 0.0), (0, 0.0)):
    pts = x[octant == r].clone()
    pts[:, 1] += shift
    out = torch.empty(pts.shape[0], device=dev)
    idx = KnnIndex(pts)
    idx.knn_self(100, out=out); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        idx.knn_self(100, out=out)
    e1.record()
    clk = []
    while not e1.query():
        clk.append(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
    torch.cuda.synchronize()
    print(json.dumps({"octant": r, "y_shift": shift, "ms": e0.elapsed_time(e1) / 3, "sm_mhz_min": min(clk) if clk else None,
                      "sm_mhz_max": max(clk) if clk else None, "power_w": pynvml.nvmlDeviceGetPowerUsage(h) / 1000}), flush=True)
    idx.close(); del pts, out
