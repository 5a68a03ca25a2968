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

from conftest import run_make

ROOT = Path(__file__).resolve().parents[1]
pytestmark = pytest.mark.skipif(platform.machine() != "x86_64", reason="the emulation's fiber switch is x86-64 assembly")


@pytest.fixture(scope="module")
def built():
    out = run_make("simt")
    assert out.returncode == 0, out.stdout[-3000:]


@pytest.mark.parametrize("world,cases", [(2, ["chunks", "tiny", "hugek"]), (3, ["radius", "clustered", "manyreturns"]), (8, ["uniform"]),
                                         (16, ["tiny"])])      # 16 = B200KNN_MAX_RANKS, more than any box has GPUs
def test_emulated_ranks_every_mode_equals_oracle(built, world, cases):
    out = subprocess.run([sys.executable, str(ROOT / "tests" / "simt" / "multirank_worker.py"), str(world), *cases],
                         cwd=str(ROOT), stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    assert out.returncode == 0 and "MULTIRANK_OK" in out.stdout, out.stdout[-4000:]


@pytest.mark.parametrize("env,expect", [({}, "one"), ({"B200KNN_SEQ_EXCHANGE": "1"}, "sequential"),
                                        ({"B200KNN_REQ_CAP": "1"}, "sequential")])
def test_emulated_one_round_exchange_and_its_fallbacks(built, env, expect):
    """The culled exchange normally takes ONE round per chunk (requests out, candidates back, merged at home);
    it falls back to world - 1 sequential rounds when switched off or when a chunk's requests do not fit their
    area. All three must give the oracle's bits, and the report must show which one ran."""
    import os
    out = subprocess.run([sys.executable, str(ROOT / "tests" / "simt" / "multirank_worker.py"), "3", "radius", "clustered"],
                         cwd=str(ROOT), env=dict(os.environ, B200KNN_EXPECT_ROUNDS=expect, **env),
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    assert out.returncode == 0 and "MULTIRANK_OK" in out.stdout, out.stdout[-4000:]


def test_emulated_ranks_are_race_free_under_thread_sanitizer(built):
    """Race detection for the protocols BETWEEN ranks (ring flags and credits, device barriers, device
    collectives, peer stores ordered by host barriers): the emulated library is compiled with
    -fsanitize=thread, every rank is a thread, and ThreadSanitizer must stay silent. (`make simt-tsan`;
    that the detector does see a missing release/acquire was checked with -DB200KNN_SIMT_BROKEN_FLAGS:
    16 reports.)"""
    import os
    try:
        tsan = subprocess.run(["/usr/bin/gcc", "-print-file-name=libtsan.so"], stdout=subprocess.PIPE, text=True).stdout.strip()
    except OSError:
        tsan = ""
    if not tsan or not os.path.isabs(tsan) or not os.path.exists(tsan) or not os.path.exists("/usr/bin/g++"):
        pytest.skip("no ThreadSanitizer runtime / system g++ on this machine")
    out = run_make("simt-tsan")
    assert out.returncode == 0, out.stdout[-3000:]
    env = dict(os.environ, LD_PRELOAD=tsan, B200KNN_SIMT_LIB=str(ROOT / "build" / "libb200knn_simt_tsan.so"),
               TSAN_OPTIONS="halt_on_error=0 report_signal_unsafe=0 history_size=4 exitcode=0")
    run = subprocess.run([sys.executable, str(ROOT / "tests" / "simt" / "multirank_worker.py"), "3", "chunks"],
                         cwd=str(ROOT), env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1500)
    assert "MULTIRANK_OK" in run.stdout, run.stdout[-4000:]
    assert "WARNING: ThreadSanitizer" not in run.stdout, run.stdout[:6000]
