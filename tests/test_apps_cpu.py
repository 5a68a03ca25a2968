"""CPU-only: the two executables + knnrun launcher, with the host back end (B200KNN_BACKEND=cpu).
Checks the reference's command-line contract, file formats and output naming, and
the multi-process bootstrap (unix-socket allgather/barrier) with 1..4 ranks."""
import os
import subprocess
from pathlib import Path

import numpy as np
import pytest

from conftest import run_make

from mpi_cuda_largescaleknn_b200.utils import generators as gen
from mpi_cuda_largescaleknn_b200.utils import io as kio
from mpi_cuda_largescaleknn_b200.utils.oracle import bit_equal, bruteforce

ROOT = Path(__file__).resolve().parents[1]
BUILD = ROOT / "build"
UV = str(BUILD / "cudaMpiKNN_unorderedData")
PV = str(BUILD / "cudaMpiKNN_prePartitionedData")
RUN = str(BUILD / "knnrun")


@pytest.fixture(scope="module", autouse=True)
def _built():
    if not (Path(UV).exists() and Path(PV).exists() and Path(RUN).exists()):
        out = run_make("all")
        assert out.returncode == 0, out.stdout[-3000:]


def run(cmd, **kw):
    env = dict(os.environ, B200KNN_BACKEND="cpu")
    return subprocess.run(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=300, **kw)


@pytest.mark.parametrize("ranks", [1, 2, 3, 4])
def test_unordered_single_output_file(tmp_path, ranks):
    n, k = 5003, 10
    x = gen.uniform(n, seed=ranks)
    inp, out = tmp_path / "in.float3", tmp_path / "out.float"
    kio.write_float3(str(inp), x)
    r = run([RUN, "-n", str(ranks), UV, str(inp), "-o", str(out), "-k", str(k)])
    assert r.returncode == 0, r.stderr
    got = kio.read_floats(str(out))
    assert got.shape == (n,)
    assert bit_equal(got, bruteforce(x, k=k))
    assert f"#0/{ranks}: got {n * 1 // ranks} points to work on" in r.stdout
    assert "done all queries..." in r.stdout


def test_unordered_without_launcher_and_with_radius(tmp_path):
    x = gen.uniform(2000, seed=5)
    inp, out = tmp_path / "in.float3", tmp_path / "o.float"
    kio.write_float3(str(inp), x)
    r = run([UV, "-k", "8", "-r", "0.05", str(inp), "-o", str(out)])      # flag order is free
    assert r.returncode == 0, r.stderr
    assert bit_equal(kio.read_floats(str(out)), bruteforce(x, k=8, max_radius=0.05))


@pytest.mark.parametrize("ranks", [2, 3])
def test_prepartitioned_per_rank_files(tmp_path, ranks):
    parts = gen.gaussian_blobs(1500, ranks, overlap=0.1, seed=2)
    names = []
    for i, p in enumerate(parts):
        f = tmp_path / f"p{i}.float3"
        kio.write_float3(str(f), p)
        names.append(str(f))
    lst = tmp_path / "files.txt"
    lst.write_text("\n".join(names))          # no trailing newline: accepted (documented divergence)
    prefix = tmp_path / "dist"
    r = run([RUN, "-n", str(ranks), PV, str(lst), "-k", "12", "-o", str(prefix)])
    assert r.returncode == 0, r.stderr
    ref = bruteforce(np.concatenate(parts), k=12)
    off = 0
    for i, p in enumerate(parts):
        name = "%s_%06d.float" % (prefix, i)
        assert os.path.exists(name)
        assert bit_equal(kio.read_floats(name), ref[off:off + len(p)])
        off += len(p)
    assert "bounds is (" in r.stdout


def test_prepartitioned_rank_count_must_match_file_count(tmp_path):
    f = tmp_path / "a.float3"
    kio.write_float3(str(f), gen.uniform(10, seed=1))
    lst = tmp_path / "files.txt"
    lst.write_text(f"{f}\n{f}\n")
    r = run([RUN, "-n", "3", PV, str(lst), "-k", "2", "-o", str(tmp_path / "x")])
    assert r.returncode != 0
    assert "does not match" in r.stderr


@pytest.mark.parametrize("argv,msg", [
    (["in.float3", "-o", "o", "-k", "0"], "invalid k"),
    (["in.float3", "-o", "o"], "no k specified"),
    (["in.float3", "-k", "3"], "no output"),
    (["-o", "o", "-k", "3"], "no input"),
    (["in.float3", "-o", "o", "-k", "3", "-z"], "unknown cmdline arg '-z'"),
])
def test_bad_command_lines_print_usage_and_exit_1(argv, msg):
    for exe in (UV, PV):
        r = run([exe] + argv)
        assert r.returncode == 1
        assert "Error:" in r.stderr and msg in r.stderr


def test_missing_input_file_is_an_empty_input(tmp_path):
    out = tmp_path / "o.float"
    r = run([UV, str(tmp_path / "nope.float3"), "-o", str(out), "-k", "3"])
    assert r.returncode == 0
    assert kio.read_floats(str(out)).shape == (0,)
