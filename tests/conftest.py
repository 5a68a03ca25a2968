import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def run_make(*targets):
    """`make -C <repo> targets...`, one at a time across all test processes (the CPU tier runs its files on several
    pytest-xdist workers; two makes that share object files must not overlap)."""
    import fcntl
    import subprocess
    (ROOT / "build").mkdir(exist_ok=True)
    with open(ROOT / "build" / ".make.lock", "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        return subprocess.run(["make", "-C", str(ROOT), *targets], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box via gpurun)")
    config.addinivalue_line("markers", "multigpu: needs >= 2 CUDA devices")
    # The CPU tier (`-m "not gpu"`) is dominated by two files that run the CUDA kernels under the host emulation;
    # give every test FILE to one of four workers (module fixtures stay per file, fixed ports stay serial inside a
    # file). Only when the caller did not choose a distribution, never for GPU runs, never inside a worker.
    # B200KNN_TEST_WORKERS=0 switches it off.
    if hasattr(config, "workerinput") or config.getoption("markexpr", "") != "not gpu":
        return
    workers = int(os.environ.get("B200KNN_TEST_WORKERS", "4"))
    if workers <= 1 or not config.pluginmanager.hasplugin("xdist"):
        return
    if getattr(config.option, "numprocesses", None) or getattr(config.option, "dist", "no") != "no":
        return
    if config.getoption("collectonly", False) or config.getoption("usepdb", False):
        return
    config.option.numprocesses = workers
    config.option.dist = "loadfile"
    config.option.tx = ["popen"] * workers


def _gpu_count():
    try:
        import torch
        return torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:
        return 0


def pytest_collection_modifyitems(config, items):
    n = _gpu_count()
    for item in items:
        if "gpu" in item.keywords and n == 0:
            item.add_marker(pytest.mark.skip(reason="no CUDA device"))
        if "multigpu" in item.keywords and n < 2:
            item.add_marker(pytest.mark.skip(reason="needs >= 2 CUDA devices"))


@pytest.fixture(scope="session")
def knnlib():
    from mpi_cuda_largescaleknn_b200.ops._lib import lib
    return lib()
