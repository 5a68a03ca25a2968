#!/usr/bin/env python
"""Throughput sweep over N for fcp, kNN and kNN with a cut-off radius on one GPU — the counterpart of
the reference library's measure.sh (submodules/cudaKDTree/measure.sh: N = 1e3..1e9, -nr 100).

    python tools/measure.py [--max-n 100000000] [--k 100] [--reps 5]   -> one JSON line per (N, op)
"""
import argparse, json, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from mpi_cuda_largescaleknn_b200.ops.knn import KnnIndex, fcp


def timed(fn, reps):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-n", type=int, default=100_000_000)
    ap.add_argument("--k", type=int, default=100)
    ap.add_argument("--reps", type=int, default=5)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    n = 1000
    while n <= a.max_n:
        x = torch.rand((n, 3), device=dev)
        out = torch.empty(n, device=dev)
        build_ms = timed(lambda: KnnIndex(x).close(), max(1, a.reps // 2))
        idx = KnnIndex(x)
        rec = {"n": n, "build_ms": build_ms,
               "fcp_ms": timed(lambda: fcp(idx, x), a.reps),
               "knn_ms": timed(lambda: idx.knn_self(a.k, out=out), a.reps),
               "knn_r0.01_ms": timed(lambda: idx.knn_self(a.k, 0.01, out=out), a.reps), "k": a.k}
        rec["knn_queries_per_s"] = n / rec["knn_ms"] * 1e3
        rec["fcp_queries_per_s"] = n / rec["fcp_ms"] * 1e3
        print(json.dumps(rec), flush=True)
        idx.close()
        del x, out
        n *= 10


if __name__ == "__main__":
    main()
