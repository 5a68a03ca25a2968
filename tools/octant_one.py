#!/usr/bin/env python
"""One octant of the bench input, two all-kNN passes (warm-up + measured) — target for ncu."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from mpi_cuda_largescaleknn_b200.ops.knn import KnnIndex
from mpi_cuda_largescaleknn_b200.utils.generators import torch_uniform
dev = torch.device("cuda", 0)
N, world, r = 256 * 1024 * 1024, 8, int(sys.argv[1])
x = torch.cat([torch_uniform(N * (i + 1) // world - N * i // world, dev, seed=1234 + i) for i in range(world)])
mid = (x.min(0).values + x.max(0).values) / 2
octant = ((x[:, 0] >= mid[0]).long() << 2) | ((x[:, 1] >= mid[1]).long() << 1) | (x[:, 2] >= mid[2]).long()
pts = x[octant == r].contiguous(); del x, octant
out = torch.empty(pts.shape[0], device=dev)
idx = KnnIndex(pts)
for _ in range(2):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); idx.knn_self(100, out=out); e1.record(); torch.cuda.synchronize()
print("octant", r, "n", pts.shape[0], "ms", e0.elapsed_time(e1))
