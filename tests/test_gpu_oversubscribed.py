"""Multi-rank paths on a single GPU: 2 and 3 ranks share device 0 (see tests/mp_onegpu_worker.py), so the
fused ring, the culled exchange, the redistribution and the device collectives are exercised — against
the brute-force oracle — wherever `pytest -m gpu` has one GPU, instead of being skipped there."""
import os
import subprocess
import sys
from pathlib import Path

import pytest

torch = pytest.importorskip("torch")

ROOT = Path(__file__).resolve().parents[1]
pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("nproc", [2, 3])
def test_ranks_sharing_one_gpu_match_oracle(nproc):
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", CUDA_VISIBLE_DEVICES=os.environ.get("CUDA_VISIBLE_DEVICES", "0").split(",")[0])
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(29580 + nproc), str(ROOT / "tests" / "mp_onegpu_worker.py")]
    out = subprocess.run(cmd, cwd=str(ROOT), env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                         timeout=600)
    assert out.returncode == 0 and "MP_ONEGPU_OK" in out.stdout, out.stdout[-4000:]


def test_unordered_executable_three_ranks_on_one_gpu(tmp_path):
    """knnrun -n 3 ... -g 1: three processes share device 0 (what the reference's -g flag is for) and each writes its
    slice of the one output file straight from its pinned result buffers; the file must hold the oracle's bits."""
    import numpy as np

    from mpi_cuda_largescaleknn_b200.utils import generators as gen
    from mpi_cuda_largescaleknn_b200.utils import io as kio
    from mpi_cuda_largescaleknn_b200.utils.oracle import bit_equal, bruteforce
    n, k = 200_003, 40
    x = gen.uniform(n, seed=23)
    inp, out = tmp_path / "in.float3", tmp_path / "out.float"
    kio.write_float3(str(inp), x)
    env = {k_: v for k_, v in os.environ.items() if k_ != "B200KNN_BACKEND"}
    r = subprocess.run([str(ROOT / "build" / "knnrun"), "-n", "3", str(ROOT / "build" / "cudaMpiKNN_unorderedData"), str(inp),
                        "-o", str(out), "-k", str(k), "-g", "1"], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                       text=True, timeout=600)
    assert r.returncode == 0, r.stderr + r.stdout
    got = kio.read_floats(str(out))
    assert got.shape == (n,)
    sel = np.arange(0, n, 97)
    assert bit_equal(got[sel], bruteforce(x, queries=x[sel], k=k))
    assert r.stdout.count("done all queries...") == 1
