"""Offline model of the insertion phase of the query kernel (no GPU needed).

The query kernel is issue bound and spends more than half of its instructions in the per-lane heap
sift loop at about 13 of 32 lanes active (profiles/README.md). This tool replays real candidate
streams — written by the CPU mirror engine with B200KNN_CPU_TRACE, which visits the same leaves in
the same order as the kernel — through per-lane 4-ary heaps and charges warp-level cost the way a
SIMT machine does: a lock-step round costs as much as its slowest lane. It compares insertion
schedules before anybody writes CUDA for them:

  per_leaf      what the kernel does: after each 32-candidate leaf, lanes loop over their survivors
  leaf_pairs    survivors of two consecutive leaves are processed together
  deferred_B    survivors are parked in a per-lane queue (B entries) and sifted when a queue is full
  fill          the first floor(k/32) leaves are appended without sifting, then one heapify

Cost unit: instructions issued by the warp (C_ROUND per lock-step round, C_LEVEL per sift level,
C_PUSH per queue append). "ideal" is the same work at 32 of 32 lanes.

    python tools/sift_model.py --n 200000 --k 100
"""
from __future__ import annotations

import argparse
import os
import subprocess
import sys
import tempfile
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

C_ROUND, C_LEVEL, C_PUSH = 14, 13, 6
REC = 8 + 8 + 32 * 4 + 32 * 32 * 4


def record_trace(n: int, k: int, path: str, kind: str) -> None:
    """Run the mirror engine in a child process (the trace file is opened once per process)."""
    code = (
        "import numpy as np, ctypes as C\n"
        "from mpi_cuda_largescaleknn_b200.ops._lib import lib\n"
        "from mpi_cuda_largescaleknn_b200.utils import generators as gen\n"
        f"x = gen.{'uniform' if kind == 'uniform' else 'clustered_surface'}({n}, seed=3)\n"
        f"out = np.empty({n}, np.float32)\n"
        f"rc = lib().b200knn_cpu_knn(x.ctypes.data, {n}, None, 0, {k}, float('inf'), out.ctypes.data, 1)\n"
        "assert rc == 0\n")
    env = dict(os.environ, B200KNN_CPU_TRACE=path)
    subprocess.run([sys.executable, "-c", code], check=True, cwd=str(ROOT), env=env)


def load_groups(path: str):
    raw = np.fromfile(path, dtype=np.uint8)
    nrec = len(raw) // REC
    raw = raw[:nrec * REC].reshape(nrec, REC)
    gid = raw[:, :8].copy().view(np.int64).ravel()
    tops = raw[:, 16:16 + 128].copy().view(np.float32).reshape(nrec, 32)
    d2 = raw[:, 16 + 128:].copy().view(np.float32).reshape(nrec, 32, 32)
    groups = []
    for g in np.unique(gid):
        sel = np.flatnonzero(gid == g)
        groups.append((tops[sel], d2[sel]))
    return groups


class Heap:
    """4-ary max-heap of k floats; replace_top returns the number of levels the sift examined."""

    def __init__(self, k: int):
        self.k = k
        self.a = [float("inf")] * k

    def top(self) -> float:
        return self.a[0]

    def replace_top(self, v: float) -> int:
        a, k, pos, levels = self.a, self.k, 0, 0
        while True:
            c = 4 * pos + 1
            if c >= k:
                break
            levels += 1
            best = c
            for j in range(c + 1, min(c + 4, k)):
                if a[j] > a[best]:
                    best = j
            if a[best] > v:
                a[pos] = a[best]
                pos = best
            else:
                break
        a[pos] = v
        return levels

    def heapify_cost(self) -> int:
        """Floyd build-heap; returns sift levels examined (convergent on a GPU: fixed loop shape)."""
        a, k, levels = self.a, self.k, 0
        for start in range((k - 2) // 4, -1, -1):
            v, pos = a[start], start
            while True:
                c = 4 * pos + 1
                if c >= k:
                    break
                levels += 1
                best = c
                for j in range(c + 1, min(c + 4, k)):
                    if a[j] > a[best]:
                        best = j
                if a[best] > v:
                    a[pos] = a[best]
                    pos = best
                else:
                    break
            a[pos] = v
        return levels


def lockstep(heaps, pending, stats) -> None:
    """pending[l] = list of candidate values of lane l; lanes advance one candidate per round."""
    rounds = max((len(p) for p in pending), default=0)
    for r in range(rounds):
        deepest, active = 0, 0
        for l, p in enumerate(pending):
            if r >= len(p):
                continue
            active += 1
            v = p[r]
            if v < heaps[l].top():
                lv = heaps[l].replace_top(v)
                stats["inserts"] += 1
                stats["lane_levels"] += lv
                deepest = max(deepest, lv)
        stats["rounds"] += 1
        stats["lane_rounds"] += active
        stats["cost"] += C_ROUND + C_LEVEL * deepest
        stats["warp_levels"] += deepest


def new_stats():
    return dict(cost=0, rounds=0, lane_rounds=0, warp_levels=0, lane_levels=0, inserts=0, pushes=0, extra=0)


def run_per_leaf(group, k, merge=1, fill=False):
    tops, d2 = group
    heaps = [Heap(k) for _ in range(32)]
    st = new_stats()
    nleaf = len(d2)
    i = 0
    if fill:
        m = k // 32
        if m > 0 and nleaf >= m:
            for l in range(32):
                vals = [float(v) for leaf in range(m) for v in d2[leaf, l] if tops[leaf, l] > -np.inf]
                h = heaps[l]
                if vals:
                    h.a[:len(vals)] = vals[:k]
            lv = max(h.heapify_cost() for h in heaps)
            st["cost"] += 32 * m * 2 + C_LEVEL * lv      # convergent appends + one heapify
            st["extra"] += 32 * m * 2 + C_LEVEL * lv
            st["inserts"] += 32 * m * 32
            i = m
    while i < nleaf:
        j = min(nleaf, i + merge)
        pending = []
        for l in range(32):
            t0 = heaps[l].top()
            if tops[i, l] == -np.inf:          # lane sits this pass out
                pending.append([])
                continue
            pending.append([float(v) for leaf in range(i, j) for v in d2[leaf, l] if v < t0])
        lockstep(heaps, pending, st)
        i = j
    return st, [h.top() for h in heaps]


def run_deferred(group, k, cap):
    tops, d2 = group
    heaps = [Heap(k) for _ in range(32)]
    queues = [[] for _ in range(32)]
    st = new_stats()

    def drain():
        lockstep(heaps, queues, st)
        for q in queues:
            q.clear()

    for i in range(len(d2)):
        if max(len(q) for q in queues) > cap - 32:
            drain()
        most = 0
        for l in range(32):
            if tops[i, l] == -np.inf:
                continue
            t0 = heaps[l].top()
            new = [float(v) for v in d2[i, l] if v < t0]
            queues[l].extend(new)
            most = max(most, len(new))
        st["pushes"] += most
        st["cost"] += C_PUSH * most
    drain()
    return st, [h.top() for h in heaps]


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=200_000)
    ap.add_argument("--k", type=int, default=100)
    ap.add_argument("--kind", default="uniform", choices=["uniform", "clustered"])
    ap.add_argument("--max-groups", type=int, default=48)
    args = ap.parse_args()
    with tempfile.TemporaryDirectory(dir=str(ROOT / "build")) as tmp:
        path = os.path.join(tmp, "trace.bin")
        record_trace(args.n, args.k, path, args.kind)
        groups = load_groups(path)[:args.max_groups]
    print(f"{len(groups)} groups, {sum(len(g[1]) for g in groups) / len(groups):.1f} leaves per group, k={args.k}")
    schemes = [("per_leaf", lambda g: run_per_leaf(g, args.k)),
               ("per_leaf+fill", lambda g: run_per_leaf(g, args.k, fill=True)),
               ("fill+pairs", lambda g: run_per_leaf(g, args.k, merge=2, fill=True)),
               ("fill+quads", lambda g: run_per_leaf(g, args.k, merge=4, fill=True)),
               ("leaf_pairs", lambda g: run_per_leaf(g, args.k, merge=2)),
               ("leaf_quads", lambda g: run_per_leaf(g, args.k, merge=4)),
               ("deferred_48", lambda g: run_deferred(g, args.k, 48)),
               ("deferred_64", lambda g: run_deferred(g, args.k, 64)),
               ("deferred_96", lambda g: run_deferred(g, args.k, 96))]
    ref_tops = None
    print(f"{'scheme':16s} {'cost/group':>11s} {'vs now':>7s} {'inserts/q':>10s} {'lanes/round':>12s} {'depth util':>11s} {'ideal':>9s}")
    base = None
    for name, fn in schemes:
        tot = new_stats()
        tops_all = []
        for g in groups:
            st, tp = fn(g)
            for key in tot:
                tot[key] += st[key]
            tops_all.append(tp)
        if ref_tops is None:
            ref_tops = tops_all
        assert tops_all == ref_tops, f"{name}: different k-th distances - the schedule is not exact"
        ng = len(groups)
        cost = tot["cost"] / ng
        base = base or cost
        ideal = ((tot["lane_rounds"] * C_ROUND + tot["lane_levels"] * C_LEVEL) / 32 + tot["extra"]) / ng
        print(f"{name:16s} {cost:11.0f} {cost / base:7.2f} {tot['inserts'] / ng / 32:10.1f} "
              f"{tot['lane_rounds'] / max(1, tot['rounds']):12.1f} "
              f"{tot['lane_levels'] / max(1, tot['warp_levels']) / max(1e-9, tot['lane_rounds'] / max(1, tot['rounds'])):11.2f} "
              f"{ideal:9.0f}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
