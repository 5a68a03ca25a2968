#!/usr/bin/env python
"""Build an index over 1 M random points and time find-closest-point queries — the counterpart of
the reference library's sample (submodules/cudaKDTree/sample.cu:80-161)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from mpi_cuda_largescaleknn_b200.ops.knn import KnnIndex, fcp
from mpi_cuda_largescaleknn_b200.utils.pretty import current_time, pretty_number, pretty_seconds

n_points, n_queries, reps = 1_000_000, 10_000_000, 10
dev = torch.device("cuda", 0)
points = torch.rand((n_points, 3), device=dev)
queries = torch.rand((n_queries, 3), device=dev)
t0 = current_time()
index = KnnIndex(points)
torch.cuda.synchronize()
print(f"built index over {pretty_number(n_points)} points in {pretty_seconds(current_time() - t0)}")
fcp(index, queries)                     # warm-up
torch.cuda.synchronize()
t0 = current_time()
for _ in range(reps):
    ids, dist = fcp(index, queries)
torch.cuda.synchronize()
dt = (current_time() - t0) / reps
print(f"{pretty_number(n_queries)} fcp queries in {pretty_seconds(dt)} = {pretty_number(n_queries / dt)} queries/s")
print("first results:", ids[:4].tolist(), dist[:4].tolist())
index.close()
