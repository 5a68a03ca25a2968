#!/usr/bin/env python
"""Time build vs query on one GPU and print traversal counters (used plain and under ncu)."""
import argparse, json, sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from mpi_cuda_largescaleknn_b200.ops.knn import KnnIndex
from mpi_cuda_largescaleknn_b200.utils.generators import torch_uniform

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=4_000_000)
ap.add_argument("--k", type=int, default=100)
ap.add_argument("--reps", type=int, default=2)
ap.add_argument("--stats", action="store_true")
ap.add_argument("--kind", default="uniform", choices=["uniform", "clustered"])
a = ap.parse_args()
dev = torch.device("cuda", 0)
if a.kind == "uniform":
    x = torch_uniform(a.n, dev, seed=1)
else:
    from tools.run_configs import clustered_surface
    x = clustered_surface(a.n, dev, 300)
out = torch.empty(a.n, device=dev)
ev = lambda: torch.cuda.Event(enable_timing=True)
for rep in range(a.reps):
    e0, e1, e2 = ev(), ev(), ev()
    e0.record()
    idx = KnnIndex(x)
    e1.record()
    if a.stats:
        idx.enable_stats(True)
    idx.knn_self(a.k, out=out)
    e2.record()
    torch.cuda.synchronize()
    rec = {"n": a.n, "k": a.k, "build_ms": e0.elapsed_time(e1), "query_ms": e1.elapsed_time(e2),
           "Mq_per_s": a.n / e1.elapsed_time(e2) / 1e3}
    if a.stats:
        st = idx.stats()
        g = (a.n + 31) // 32
        rec.update({"leaves_per_group": st["leaves_visited"] / g, "inner_per_group": st["inner_nodes"] / g,
                    "inserts_per_query": st["heap_inserts"] / a.n,
                    "fallbacks_per_group": st["dfs_fallbacks"] / g, "purges_per_group": st["purges"] / g})
    print(json.dumps(rec), flush=True)
    idx.close()
