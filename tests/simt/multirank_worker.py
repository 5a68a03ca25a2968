"""Several emulated GPUs in one process: one Python thread per rank calls the C ABI of the host-emulated
library (tests/simt); ctypes releases the GIL, so the ranks' kernels really run concurrently, store into
each other's "peer" memory and hand-shake through the same sequence flags as on NVLink. The bootstrap
(allgather + barrier of host bytes) is a threading.Barrier.

    python tests/simt/multirank_worker.py <world> [case ...]      prints MULTIRANK_OK or the failures
"""
from __future__ import annotations

import ctypes as C
import math
import os
import sys
import threading
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

from mpi_cuda_largescaleknn_b200.ops import _lib  # noqa: E402
from mpi_cuda_largescaleknn_b200.parallel.dist import ALLGATHER_FN, BARRIER_FN, Bootstrap, DistReport, MODES  # noqa: E402
from mpi_cuda_largescaleknn_b200.utils import generators as gen  # noqa: E402
from mpi_cuda_largescaleknn_b200.utils.oracle import bit_equal, bruteforce  # noqa: E402


def load():
    import os
    lib = _lib._proto(C.CDLL(os.environ.get("B200KNN_SIMT_LIB", str(ROOT / "build" / "libb200knn_simt.so"))))
    lib.b200knn_dist_create.argtypes = [C.POINTER(Bootstrap), C.POINTER(C.c_void_p)]
    lib.b200knn_dist_knn.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int, C.c_float, C.c_int, C.c_int,
                                     C.c_void_p, C.c_void_p, C.POINTER(DistReport)]
    lib.b200knn_dist_destroy.argtypes = [C.c_void_p]
    lib.b200knn_dist_allgather_dev.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    lib.b200knn_dist_allreduce_dev.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    return lib


class World:
    """The bootstrap every rank hands to the library."""

    def __init__(self, n):
        self.n = n
        self.slots = [b""] * n
        self.bar = threading.Barrier(n, timeout=240)

    def bootstrap(self, rank):
        def allgather(ctx, send, recv, nbytes):
            try:
                self.slots[rank] = C.string_at(send, nbytes)
                self.bar.wait()
                data = b"".join(self.slots)
                C.memmove(recv, data, len(data))
                self.bar.wait()
                return 0
            except Exception as e:  # noqa: BLE001
                print(f"[rank {rank}] allgather failed: {e!r}", flush=True)
                return 1

        def barrier(ctx):
            try:
                self.bar.wait()
                return 0
            except Exception as e:  # noqa: BLE001
                print(f"[rank {rank}] barrier failed: {e!r}", flush=True)
                return 1

        ag, br = ALLGATHER_FN(allgather), BARRIER_FN(barrier)
        return Bootstrap(None, rank, self.n, ag, br), (ag, br)


def main() -> int:
    world = int(sys.argv[1])
    wanted = set(sys.argv[2:])
    lib = load()
    w = World(world)
    failures = []
    lock = threading.Lock()

    def fail(msg):
        with lock:
            failures.append(msg)
            print(msg, flush=True)

    # (name, total points, k, max radius, groups per chunk, generator)
    cases = [("uniform", 2400 * world, 16, math.inf, 0, "uniform"),
             ("chunks", 1500 * world + 7, 40, math.inf, 13, "uniform"),       # several chunks, fill path, ragged split
             ("radius", 1800 * world, 32, 0.08, 0, "uniform"),
             ("clustered", 2000 * world, 10, math.inf, 0, "clustered"),
             ("tiny", world * 3 + 1, 5, math.inf, 0, "uniform"),
             ("hugek", 1300 * world, 1900, math.inf, 0, "uniform"),     # heaps in global memory, every group travels
             ("manyreturns", 3000 * world + 1, 100, math.inf, 16, "uniform")]   # unpartitioned: every group asks every rank
    cases = [c for c in cases if (c[0] in wanted if wanted else c[0] not in ("hugek", "manyreturns"))]

    def rank_main(rank):
        bs, keep = w.bootstrap(rank)
        eng = C.c_void_p()
        if lib.b200knn_dist_create(C.byref(bs), C.byref(eng)):
            return fail(f"[rank {rank}] dist_create: {lib.b200knn_last_error().decode()}")
        # device-side collectives against numpy
        rng = np.random.default_rng(99)                       # same stream on every rank
        for it in range(30):
            words = int(rng.integers(1, 33))
            table = rng.integers(-2**31, 2**31 - 1, (world, words)).astype(np.int32)
            got = np.zeros((world, words), np.int32)
            mine = np.ascontiguousarray(table[rank])
            if lib.b200knn_dist_allgather_dev(eng, mine.ctypes.data, words * 4, got.ctypes.data, None):
                return fail(f"[rank {rank}] allgather_dev: {lib.b200knn_last_error().decode()}")
            if not np.array_equal(got, table):
                fail(f"[rank {rank}] allgather_dev mismatch at iteration {it}")
            count = int(rng.integers(1, 17))
            t64 = rng.integers(-2**58, 2**58, (world, count)).astype(np.int64)
            for op, want in ((0, t64.max(0)), (1, t64.sum(0)), (2, t64.min(0))):
                out = np.zeros(count, np.int64)
                mine64 = np.ascontiguousarray(t64[rank])
                if lib.b200knn_dist_allreduce_dev(eng, mine64.ctypes.data, count, op, out.ctypes.data, None):
                    return fail(f"[rank {rank}] allreduce_dev: {lib.b200knn_last_error().decode()}")
                if not np.array_equal(out, want):
                    fail(f"[rank {rank}] allreduce_dev op {op} mismatch at iteration {it}")
        # the distributed driver, every strategy
        for name, N, k, r, gc, kind in cases:
            x_all = gen.uniform(N, seed=N) if kind == "uniform" else gen.clustered_surface(N, seed=N)
            b, e = N * rank // world, N * (rank + 1) // world
            x = np.ascontiguousarray(x_all[b:e])
            ref = bruteforce(x_all, queries=x, k=k, max_radius=r, threads=1) if len(x) else np.zeros(0, np.float32)
            for mode in ("ring", "prepartitioned", "redistribute", "auto"):
                out = np.full(len(x), -1.0, np.float32)
                rep = DistReport()
                rc = lib.b200knn_dist_knn(eng, x.ctypes.data if len(x) else None, len(x), b, k, r, MODES[mode], gc,
                                          out.ctypes.data if len(x) else None, None, C.byref(rep))
                if rc:
                    return fail(f"[rank {rank}] {name}/{mode}: rc={rc} {lib.b200knn_last_error().decode()}")
                if not bit_equal(out, ref):
                    bad = int((out.view(np.uint32) != ref.view(np.uint32)).sum())
                    fail(f"[rank {rank}] {name}/{mode}: {bad} of {len(x)} results differ from the oracle")
                # which exchange ran: one round per chunk, or world - 1 sequential rounds (B200KNN_EXPECT_ROUNDS)
                expect = os.environ.get("B200KNN_EXPECT_ROUNDS")
                if expect and mode in ("prepartitioned", "redistribute") and rep.chunks == 1:
                    want = 1 if expect == "one" else world - 1
                    if rep.rounds != want:
                        fail(f"[rank {rank}] {name}/{mode}: {rep.rounds} exchange rounds, expected {want}")
            if rank == 0:
                print(f"case {name}: N={N} k={k} r={r} gc={gc} done", flush=True)
        lib.b200knn_dist_destroy(eng)
        del keep

    threads = [threading.Thread(target=rank_main, args=(r,), daemon=True) for r in range(world)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    print("MULTIRANK_OK" if not failures else f"MULTIRANK_FAILED {len(failures)}", flush=True)
    return 0 if not failures else 1


if __name__ == "__main__":
    sys.exit(main())
