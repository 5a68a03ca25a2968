"""CPU-only: SEVERAL emulated GPUs. Each rank is a thread that drives the host-emulated library
(tests/simt): the fused ring with its in-kernel sequence flags and credits, the culled exchange with
device-side barriers and the device-side vote, redistribution by peer stores, the device collectives —
the same kernels as on NVLink, "peer memory" being ordinary shared memory. Every rank's results must equal
the brute-force oracle over the union of all ranks' points, in every mode."""
import platform
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
pytestmark = pytest.mark.skipif(platform.machine() != "x86_64", reason="the emulation's fiber switch is x86-64 assembly")


@pytest.fixture(scope="module")
def built():
    out = subprocess.run(["make", "-C", str(ROOT), "simt"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert out.returncode == 0, out.stdout[-3000:]


@pytest.mark.parametrize("world,cases", [(2, ["uniform", "chunks", "tiny"]), (3, ["chunks", "radius", "clustered"]),
                                         (8, ["uniform", "tiny"])])
def test_emulated_ranks_every_mode_equals_oracle(built, world, cases):
    out = subprocess.run([sys.executable, str(ROOT / "tests" / "simt" / "multirank_worker.py"), str(world), *cases],
                         cwd=str(ROOT), stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    assert out.returncode == 0 and "MULTIRANK_OK" in out.stdout, out.stdout[-4000:]
