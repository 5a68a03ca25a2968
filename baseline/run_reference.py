"""bench.py --impl reference: run the UNMODIFIED reference (baseline/_ref, built by build_ref.sh from
/root/reference) through its own command line for the same metric and config, and print the same
JSON line with "impl": "reference". None of this repository's kernels or engine is on that path.

How the reference is run
  * no MPI on the image -> ranks are plain processes; baseline/mpi_shim provides the ten MPI calls
    (device buffers by CUDA IPC rendezvous = one device-to-device copy over NVLink, as a CUDA-aware
    MPI would do; tiny host messages through files in /dev/shm).
  * the unmodified reference needs pointsPerRank * k < 2^31 (int overflow), so the
    number of reference ranks is the smallest multiple of the GPU count that satisfies it, spread with
    the reference's own -g flag (rank % gpus). If those ranks' candidate heaps (n*k*8 bytes each) do
    not fit the GPUs' memory the configuration cannot be run -> {"unavailable": ...}.
  * the reference has no timers: the shim records wall-clock marks. "value" uses
    [first MPI_Comm_size .. first MPI_Barrier] = file read + H2D + tree build + all rounds + extract
    (max over ranks); "e2e" is the wall time of the whole launch including CUDA start-up and output.
"""
from __future__ import annotations

import json
import math
import os
import shutil
import subprocess
import sys
import tempfile
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
REF_DIR = ROOT / "baseline" / "_ref"
EXE_UNORDERED = REF_DIR / "cudaMpiKNN_unorderedData"
EXE_PREPART = REF_DIR / "cudaMpiKNN_prePartitionedData"
INT_MAX = 2 ** 31 - 1
GPU_BYTES = 178 * 1024 ** 3


def ensure_built() -> str | None:
    if EXE_UNORDERED.exists() and EXE_PREPART.exists():
        return None
    try:
        out = subprocess.run(["bash", str(ROOT / "baseline" / "build_ref.sh")], stdout=subprocess.PIPE,
                             stderr=subprocess.STDOUT, text=True, timeout=900)
    except Exception as e:  # noqa: BLE001
        return f"reference build failed: {e!r}"
    if out.returncode != 0 or not EXE_UNORDERED.exists():
        return "reference build failed: " + out.stdout[-300:].replace("\n", " ")
    return None


def make_input(n: int, path: Path, seed: int = 1234) -> None:
    """Uniform-random float3 in the unit cube, written in chunks (3 GB at the headline size)."""
    if path.exists() and path.stat().st_size == n * 12:
        return
    rng = np.random.default_rng(seed)
    with open(path, "wb") as f:
        left = n
        while left > 0:
            m = min(left, 16 * 1024 * 1024)
            rng.random((m, 3), dtype=np.float32).tofile(f)
            left -= m


def plan_ranks(n: int, k: int, gpus: int):
    """Smallest multiple of `gpus` ranks with n_rank*k < 2^31 and n_rank*12 < 2^31; memory check."""
    need = max(math.ceil(n * k / INT_MAX), math.ceil(n * 12 / INT_MAX), 1)
    ranks = gpus * math.ceil(need / gpus)
    while math.ceil(n / ranks) * k > INT_MAX:
        ranks += gpus
    n_rank = math.ceil(n / ranks)
    per_rank = n_rank * k * 8 + n_rank * 12 * 3 + n_rank * 24 + n_rank * 4 + (600 << 20)   # heaps, 2 trees + queries, sort temps, result, context
    per_gpu = per_rank * (ranks // gpus)
    return ranks, n_rank, per_gpu


def run_once(exe: Path, in_arg: str, out_arg: str, k: int, ranks: int, gpus: int, timeout_s: int):
    shim_dir = Path(tempfile.mkdtemp(prefix="b200knn_shim_", dir="/dev/shm"))
    procs = []
    t0 = time.time()
    try:
        for r in range(ranks):
            env = dict(os.environ, SHIM_RANK=str(r), SHIM_SIZE=str(ranks), SHIM_DIR=str(shim_dir))
            for v in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
                env.pop(v, None)
            cmd = [str(exe), in_arg, "-o", out_arg, "-k", str(k), "-g", str(gpus)]
            procs.append(subprocess.Popen(cmd, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True))
        deadline = t0 + timeout_s
        errs = []
        for p in procs:
            try:
                _, err = p.communicate(timeout=max(1, deadline - time.time()))
            except subprocess.TimeoutExpired:
                for q in procs:
                    q.kill()
                return None, f"reference run exceeded {timeout_s} s"
            if p.returncode != 0:
                errs.append((p.returncode, (err or "")[-200:].replace("\n", " ")))
        wall = time.time() - t0
        if errs:
            for q in procs:
                if q.poll() is None:
                    q.kill()
            return None, f"reference rank failed: rc={errs[0][0]} {errs[0][1]}"
        marks = []
        for r in range(ranks):
            marks.append([float(x) for x in (shim_dir / f"timing_{r}").read_text().split()])
        marks = np.array(marks)
        return {"wall_s": wall, "t_first": marks[:, 0].min(), "t_first_barrier": marks[:, 1].max(),
                "t_last_barrier": marks[:, 2].max(), "t_final": marks[:, 3].max()}, None
    finally:
        shutil.rmtree(shim_dir, ignore_errors=True)


def bench_reference(args, variant: str = "unmodified") -> int:
    """variant "unmodified": the reference as it is (what `--impl reference` runs, oversubscribed if n*k needs it).
    variant "i64": the three-expression size_t patch (baseline/ref_i64.patch) at exactly one rank per GPU."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0      # the reference launch is driven by one process; torchrun's other ranks have nothing to do

    def emit(d):
        print(json.dumps(d), flush=True)
        return 0

    n, k, gpus = int(args.n), int(args.k), int(args.gpus)
    err = ensure_built()
    if err:
        return emit({"impl": "reference", "unavailable": err})
    exe = EXE_UNORDERED
    ranks, n_rank, per_gpu = plan_ranks(n, k, gpus)
    if variant == "i64":
        exe = ROOT / "baseline" / "_ref" / "cudaMpiKNN_unorderedData_i64"
        if not exe.exists():
            out = subprocess.run(["bash", str(ROOT / "baseline" / "build_ref_i64.sh")], stdout=subprocess.PIPE,
                                 stderr=subprocess.STDOUT, text=True, timeout=900)
            if out.returncode != 0 or not exe.exists():
                return emit({"impl": "reference-i64", "unavailable": "build failed: " + out.stdout[-300:].replace("\n", " ")})
        ranks, n_rank = gpus, math.ceil(n / gpus)
        per_gpu = n_rank * k * 8 + n_rank * 12 * 3 + n_rank * 24 + n_rank * 4 + (600 << 20)
    if per_gpu > GPU_BYTES:
        return emit({"impl": "reference", "n_gpus": gpus,
                     "unavailable": f"unmodified reference needs {ranks} ranks (int32 limit n*k<2^31) = "
                                    f"{per_gpu / 2**30:.0f} GiB of candidate heaps+trees per GPU on {gpus} GPU(s); exceeds 180 GB"})
    work = Path("/dev/shm") if Path("/dev/shm").is_dir() else Path(tempfile.gettempdir())
    in_path = work / f"b200knn_ref_{n}.float3"
    out_path = work / f"b200knn_ref_{n}_{k}.float"
    try:
        make_input(n, in_path)
    except OSError as e:
        return emit({"impl": "reference", "unavailable": f"cannot write the {n * 12 / 2**30:.1f} GiB input file: {e}"})
    budget = int(os.environ.get("B200KNN_REF_TIMEOUT", "1500"))
    # A full reference run at the headline size takes minutes, so: run once; if that run was short
    # (< 60 s) treat it as the warm-up and time `steps` more runs, otherwise it IS the timed step.
    runs, steps_done, warm_done = [], 0, 0
    first, why = run_once(exe, str(in_path), str(out_path), k, ranks, gpus, budget)
    if first is None:
        return emit({"impl": "reference", "n_gpus": gpus, "unavailable": why})
    if first["wall_s"] < 60.0 and int(args.warmup) > 0:
        warm_done = 1
        for _ in range(max(1, int(args.steps))):
            res, why = run_once(exe, str(in_path), str(out_path), k, ranks, gpus, budget)
            if res is None:
                return emit({"impl": "reference", "n_gpus": gpus, "unavailable": why})
            runs.append(res)
            steps_done += 1
    else:
        runs.append(first)
        steps_done = 1
    dev_s = float(np.mean([r["t_first_barrier"] - r["t_first"] for r in runs]))
    wall_s = float(np.mean([r["wall_s"] for r in runs]))
    line = {
        "metric": "kNN queries/sec (N=256M float3, k=100)" if (n == 256 * 1024 * 1024 and k == 100)
                  else f"kNN queries/sec (N={n} float3, k={k})",
        "value": n / dev_s, "unit": "queries/s", "n_gpus": gpus, "steps": steps_done, "warmup": warm_done,
        "ms_per_step": dev_s * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "fp32", "data": "synthetic uniform-random float3 in the unit cube (file in /dev/shm)",
        "impl": "reference" if variant == "unmodified" else "reference-i64",
        "config": {"model": "unorderedDataVariant", "n_points": n, "k": k, "global_batch": n, "seq_len": k,
                   "parallelism": (f"mpi-shim ranks={ranks} (-g {gpus}; {ranks // gpus} per GPU, int32 limit n*k<2^31)"
                                   if variant == "unmodified" else
                                   f"mpi-shim ranks={ranks} (one per GPU; reference with baseline/ref_i64.patch)"),
                   "step": "file read + H2D + cukd::buildTree + all rounds of runQuery + extractFinalResult "
                           "(first MPI_Comm_size .. first MPI_Barrier, wall clock, max over ranks)",
                   "timing": "host wall clock marks in the MPI shim; the reference has no device timers"},
        "e2e": {"value": n / wall_s, "unit": "queries/s", "ms_per_step": wall_s * 1e3,
                "h2d_bytes_per_step": int(n * 24), "d2h_bytes_per_step": int(n * 4),
                "note": "whole launch: process start, CUDA init, file read, compute, output file"},
        "gpu_launches": None,
    }
    try:
        os.unlink(out_path)
    except OSError:
        pass
    return emit(line)
