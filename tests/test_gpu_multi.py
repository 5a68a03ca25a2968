"""Multi-GPU tests: launched through torchrun, one rank per GPU, NCCL for plumbing only."""
import os
import subprocess
import sys
from pathlib import Path

import pytest

torch = pytest.importorskip("torch")

ROOT = Path(__file__).resolve().parents[1]
pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _torchrun(script: str, nproc: int, timeout: int = 900) -> subprocess.CompletedProcess:
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", "29571", str(ROOT / "tests" / script)]
    return subprocess.run(cmd, cwd=str(ROOT), env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                          text=True, timeout=timeout)


def test_fused_and_baseline_ring_match_oracle():
    n = min(torch.cuda.device_count(), 8)
    out = _torchrun("mp_ring_worker.py", n)
    assert out.returncode == 0 and "MP_RING_OK" in out.stdout, out.stdout[-4000:]


def test_distributed_driver_all_modes_match_oracle():
    n = min(torch.cuda.device_count(), 8)
    out = _torchrun("mp_dist_worker.py", n)
    assert out.returncode == 0 and "MP_DIST_OK" in out.stdout, out.stdout[-4000:]


def test_device_collectives_match_torch_distributed():
    n = min(torch.cuda.device_count(), 8)
    out = _torchrun("mp_coll_worker.py", n, timeout=600)
    assert out.returncode == 0 and "MP_COLL_OK" in out.stdout, out.stdout[-4000:]
