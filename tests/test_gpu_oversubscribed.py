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
