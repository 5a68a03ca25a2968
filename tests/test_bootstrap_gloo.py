"""CPU-only, multi-process (gloo): host-side plumbing of the distributed driver."""
import os
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


@pytest.mark.parametrize("world", [2, 3])
def test_torch_bootstrap_over_gloo(world):
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29610 + world), str(ROOT / "tests" / "mp_bootstrap_worker.py")]
    out = subprocess.run(cmd, cwd=str(ROOT), env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert out.returncode == 0 and "MP_BOOTSTRAP_OK" in out.stdout, out.stdout[-3000:]
