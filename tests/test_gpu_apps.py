"""GPU: the executables end to end through knnrun (one rank per visible GPU), default back end."""
import os
import subprocess
from pathlib import Path

import numpy as np
import pytest

torch = pytest.importorskip("torch")

from mpi_cuda_largescaleknn_b200.utils import generators as gen
from mpi_cuda_largescaleknn_b200.utils import io as kio
from mpi_cuda_largescaleknn_b200.utils.oracle import bit_equal, bruteforce

ROOT = Path(__file__).resolve().parents[1]
BUILD = ROOT / "build"
pytestmark = pytest.mark.gpu


def _ranks():
    return max(1, min(torch.cuda.device_count(), 8))


def _run(cmd):
    env = {k: v for k, v in os.environ.items() if k != "B200KNN_BACKEND"}
    return subprocess.run(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)


@pytest.mark.parametrize("extra", [[], ["--ring"], ["--in-place"]])
def test_unordered_executable_on_gpus(tmp_path, extra):
    n, k = 300_007, 100
    x = gen.uniform(n, seed=17)
    inp, out = tmp_path / "in.float3", tmp_path / "out.float"
    kio.write_float3(str(inp), x)
    r = _run([str(BUILD / "knnrun"), "-n", str(_ranks()), str(BUILD / "cudaMpiKNN_unorderedData"), str(inp),
              "-o", str(out), "-k", str(k)] + extra)
    assert r.returncode == 0, r.stderr + r.stdout
    got = kio.read_floats(str(out))
    assert got.shape == (n,)
    sel = np.arange(0, n, 150)
    assert bit_equal(got[sel], bruteforce(x, queries=x[sel], k=k))
    assert "timing(rank 0)" in r.stdout


def test_prepartitioned_executable_on_gpus(tmp_path):
    ranks = _ranks()
    parts = gen.gaussian_blobs(50_000, ranks, overlap=0.1, seed=4)
    names = []
    for i, p in enumerate(parts):
        f = tmp_path / f"p{i}.float3"
        kio.write_float3(str(f), p)
        names.append(str(f))
    lst = tmp_path / "files.txt"
    lst.write_text("\n".join(names) + "\n")
    prefix = tmp_path / "dist"
    r = _run([str(BUILD / "knnrun"), "-n", str(ranks), str(BUILD / "cudaMpiKNN_prePartitionedData"), str(lst),
              "-k", "32", "-o", str(prefix)])
    assert r.returncode == 0, r.stderr + r.stdout
    x_all = np.concatenate(parts)
    off = 0
    for i, p in enumerate(parts):
        got = kio.read_floats(kio.per_rank_output_name(str(prefix), i))
        sel = np.arange(0, len(p), 50)
        assert bit_equal(got[sel], bruteforce(x_all, queries=p[sel], k=32))
        off += len(p)
