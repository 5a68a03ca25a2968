"""Offline model of the load balance of the redistributed pipeline (no GPU needed).

The redistribution gives every rank one contiguous range of the global Morton order with (almost)
the same number of points. On clustered data the ranks' local phases then differ by 3.5x (measured:
153-535 ms, 512 Mi points, k = 32, 8 GPUs) because the work per query group depends on where the group
lies, not on how many points the rank owns. This tool measures that work with the CPU mirror engine
(leaves visited per rank, the quantity the kernel's time follows) for alternative ownership maps:

  contiguous     R ranges of the Morton order (what dist.cu does today)
  interleaved_c  R*c ranges dealt round-robin, c = 4, 8, 16: every rank owns c scattered pieces

and reports max/mean work (1.0 = perfectly balanced) and the share of groups whose k-th-neighbour ball
leaves the rank's own pieces (the groups that would have to visit another rank).

    python tools/balance_model.py --n 400000 --k 32 --ranks 8 --kind clustered
"""
from __future__ import annotations

import argparse
import ctypes as C
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

from mpi_cuda_largescaleknn_b200.ops._lib import lib  # noqa: E402
from mpi_cuda_largescaleknn_b200.utils import generators as gen  # noqa: E402
from mpi_cuda_largescaleknn_b200.utils.oracle import cpu_knn  # noqa: E402


def morton_order(x: np.ndarray) -> np.ndarray:
    lo, hi = x.min(0), x.max(0)
    q = np.minimum(((x - lo) / np.maximum(hi - lo, 1e-30) * 1023.0).astype(np.uint64), np.uint64(1023))

    def spread(v):
        v = (v | (v << np.uint64(16))) & np.uint64(0x030000FF)
        v = (v | (v << np.uint64(8))) & np.uint64(0x0300F00F)
        v = (v | (v << np.uint64(4))) & np.uint64(0x030C30C3)
        v = (v | (v << np.uint64(2))) & np.uint64(0x09249249)
        return v

    key = (spread(q[:, 0]) << np.uint64(2)) | (spread(q[:, 1]) << np.uint64(1)) | spread(q[:, 2])
    return np.argsort(key, kind="stable")


def rank_work(x_rank: np.ndarray, k: int) -> tuple[int, np.ndarray]:
    """Leaves visited by the local phase of a rank that owns x_rank, and its local k-th distances."""
    st = (C.c_uint64 * 4)()
    lib().b200knn_cpu_stats(st, 1)
    d = cpu_knn(x_rank, k=k)
    lib().b200knn_cpu_stats(st, 1)
    return int(st[0]), d


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=400_000)
    ap.add_argument("--k", type=int, default=32)
    ap.add_argument("--ranks", type=int, default=8)
    ap.add_argument("--kind", default="clustered", choices=["clustered", "uniform"])
    args = ap.parse_args()
    x = gen.clustered_surface(args.n, seed=5) if args.kind == "clustered" else gen.uniform(args.n, seed=5)
    order = morton_order(x)
    xs = x[order]
    d_global = cpu_knn(xs, k=args.k)            # true k-th distances (all points)
    R = args.ranks
    print(f"{args.kind}, n={args.n}, k={args.k}, ranks={R}")
    print(f"{'ownership':16s} {'max/mean work':>14s} {'min/mean':>9s} {'groups leaving home':>20s}")
    for c in (1, 4, 8, 16):
        pieces = np.array_split(np.arange(args.n), R * c)
        work, leaving = [], 0
        for r in range(R):
            own = np.concatenate([pieces[i] for i in range(r, R * c, R)])
            w, d_local = rank_work(xs[own], args.k)
            work.append(w)
            # a point whose local k-th distance is larger than the global one has neighbours elsewhere;
            # a group leaves home if any of its 32 queries does (approximation of the box test)
            worse = d_local > d_global[own]
            ng = (len(own) + 31) // 32
            pad = np.zeros(ng * 32, bool)
            pad[:len(own)] = worse
            leaving += int(pad.reshape(ng, 32).any(1).sum())
        work = np.array(work, float)
        name = "contiguous" if c == 1 else f"interleaved_{c}"
        print(f"{name:16s} {work.max() / work.mean():14.2f} {work.min() / work.mean():9.2f} "
              f"{100.0 * leaving / ((args.n + 31) // 32):19.1f}%")
    return 0


if __name__ == "__main__":
    sys.exit(main())
