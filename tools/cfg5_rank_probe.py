#!/usr/bin/env python
"""One GPU: build BASELINE config 5 (clustered/surface) at full size, cut it into the 8 Morton-order shares an
8-GPU redistribution would produce, and time a self-kNN on each share separately (local phase of each rank,
without the other ranks). Finds which share is slow and why (traversal counters)."""
import argparse, json, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from mpi_cuda_largescaleknn_b200.ops.knn import KnnIndex
from tools.run_configs import clustered_surface

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=512 * 1024 * 1024)
ap.add_argument("--ranks", type=int, default=8)
ap.add_argument("--k", type=int, default=32)
ap.add_argument("--only", type=int, default=-1, help="probe just this share")
ap.add_argument("--dump", default="", help="write the probed share's points to this .npy file")
a = ap.parse_args()
dev = torch.device("cuda", 0)
parts = [clustered_surface(a.n // a.ranks, dev, 300 + r) for r in range(a.ranks)]
x = torch.cat(parts); del parts
lo, hi = x.min(0).values, x.max(0).values
q = ((x - lo) / (hi - lo) * 2097151.0).to(torch.int64).clamp_(0, 2097151)
def spread(v):
    v = v & 0x1fffff
    v = (v | (v << 32)) & 0x1f00000000ffff
    v = (v | (v << 16)) & 0x1f0000ff0000ff
    v = (v | (v << 8)) & 0x100f00f00f00f00f
    v = (v | (v << 4)) & 0x10c30c30c30c30c3
    v = (v | (v << 2)) & 0x1249249249249249
    return v
key = (spread(q[:, 0]) << 2) | (spread(q[:, 1]) << 1) | spread(q[:, 2]); del q
order = torch.argsort(key); del key
for r in range(a.ranks):
    if a.only >= 0 and r != a.only:
        continue
    sel = order[a.n * r // a.ranks: a.n * (r + 1) // a.ranks]
    xr = x[sel].contiguous()
    out = torch.empty(xr.shape[0], device=dev)
    if a.dump:
        import numpy as np
        np.save(a.dump, xr.cpu().numpy())
    for rep in range(2):
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        e0.record(); idx = KnnIndex(xr); e1.record(); idx.enable_stats(True); idx.knn_self(a.k, out=out); e2.record()
        torch.cuda.synchronize()
        st = idx.stats(); g = (xr.shape[0] + 31) // 32
        if rep == 1 and a.only >= 0:
            import ctypes as C, numpy as np
            from mpi_cuda_largescaleknn_b200.ops._lib import lib
            srt = np.empty((xr.shape[0], 4), np.float32)
            lib().b200knn_index_copy_sorted(idx._h, srt.ctypes.data)
            gsl = st["slowest_group"]
            print("slowest group", gsl, "kcycles", st["slowest_group_kcycles"])
            for d in (-1, 0, 1):
                blk = srt[(gsl + d) * 32:(gsl + d + 1) * 32, :3]
                print("group", gsl + d, "lo", blk.min(0), "hi", blk.max(0))
            np.set_printoptions(precision=7, suppress=False, linewidth=200)
            print(srt[gsl * 32:(gsl + 1) * 32, :3])
            res = out.cpu().numpy()
            ids = srt[gsl * 32:(gsl + 1) * 32, 3].copy().view(np.uint32)
            print("kth distances", res[ids])
        idx.close()
    print(json.dumps({"share": r, "n": int(xr.shape[0]), "build_ms": e0.elapsed_time(e1), "query_ms": e1.elapsed_time(e2),
                      "leaves_per_group": st["leaves_visited"] / g, "inner_per_group": st["inner_nodes"] / g,
                      "inserts_per_query": st["heap_inserts"] / xr.shape[0], "fallbacks_per_group": st["dfs_fallbacks"] / g, "stats": st,
                      "extent": [round(float(v), 4) for v in (xr.max(0).values - xr.min(0).values)]}), flush=True)
    del xr, out
