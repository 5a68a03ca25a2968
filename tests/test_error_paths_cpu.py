"""CPU-only: the failure paths. Library code never throws or exits: it returns a non-zero code and leaves a
message in b200knn_last_error(). These tests drive the limits and argument checks through the C ABI of the
host-emulated library (tests/simt; the checks are host code and identical in the GPU build), and the launcher's
propagation of a failing rank's exit code."""
import ctypes as C
import math
import os
import platform
import subprocess
from pathlib import Path

import numpy as np
import pytest

from conftest import run_make

from mpi_cuda_largescaleknn_b200.ops import _lib
from mpi_cuda_largescaleknn_b200.utils import generators as gen
from mpi_cuda_largescaleknn_b200.utils import io as kio

ROOT = Path(__file__).resolve().parents[1]
pytestmark = pytest.mark.skipif(platform.machine() != "x86_64", reason="the emulation's fiber switch is x86-64 assembly")


@pytest.fixture(scope="module")
def sim():
    out = run_make("simt", "all")
    assert out.returncode == 0, out.stdout[-3000:]
    lib = _lib._proto(C.CDLL(os.environ.get("B200KNN_SIMT_LIB", str(ROOT / "build" / "libb200knn_simt.so"))))
    return lib


def err(lib):
    return lib.b200knn_last_error().decode()


def test_argument_checks_return_codes_and_messages(sim):
    x = gen.uniform(500, seed=1)
    h = C.c_void_p()
    assert sim.b200knn_index_create(x.ctypes.data, -1, None, 0, None, C.byref(h)) != 0
    assert "n must be in [0, 2^31-64)" in err(sim)
    sim.b200knn_index_destroy(h)
    assert sim.b200knn_index_create(x.ctypes.data, 2**31, None, 0, None, C.byref(h)) != 0      # per-GPU point limit
    assert "2^31-64" in err(sim)
    sim.b200knn_index_destroy(h)
    assert sim.b200knn_index_create(x.ctypes.data, len(x), None, 0, None, None) != 0
    assert "out is NULL" in err(sim)
    h = C.c_void_p()
    assert sim.b200knn_index_create(x.ctypes.data, len(x), None, 0, None, C.byref(h)) == 0
    out = np.zeros(len(x), np.float32)
    assert sim.b200knn_knn_self(h, 0, math.inf, out.ctypes.data, None) != 0
    assert "k must be >= 1" in err(sim)
    assert sim.b200knn_knn_self(h, 3, math.inf, None, None) != 0
    assert "NULL argument" in err(sim)
    assert sim.b200knn_knn_query(h, x.ctypes.data, 2**31, 3, math.inf, out.ctypes.data, None) != 0
    assert "nq out of range" in err(sim)
    # a good call still works afterwards: errors leave no state behind
    assert sim.b200knn_knn_self(h, 3, math.inf, out.ctypes.data, None) == 0
    assert np.all(np.isfinite(out))
    sim.b200knn_index_destroy(h)


def test_distributed_engine_limits(sim):
    class Bootstrap(C.Structure):
        _fields_ = [("ctx", C.c_void_p), ("rank", C.c_int), ("world", C.c_int), ("allgather", C.c_void_p), ("barrier", C.c_void_p)]
    sim.b200knn_dist_create.argtypes = [C.POINTER(Bootstrap), C.POINTER(C.c_void_p)]
    sim.b200knn_dist_create.restype = C.c_int
    eng = C.c_void_p()
    assert sim.b200knn_dist_create(C.byref(Bootstrap(None, 0, 17, None, None)), C.byref(eng)) != 0     # more than 16 ranks
    assert "world must be in [1,16]" in err(sim)
    assert sim.b200knn_dist_create(C.byref(Bootstrap(None, 3, 2, None, None)), C.byref(eng)) != 0
    assert "rank out of range" in err(sim)
    assert sim.b200knn_dist_create(C.byref(Bootstrap(None, 0, 2, None, None)), C.byref(eng)) != 0      # 2 ranks, no callbacks
    assert "bootstrap callbacks missing" in err(sim)
    assert sim.b200knn_dist_create(C.byref(Bootstrap(None, 0, 1, None, None)), C.byref(eng)) == 0
    sim.b200knn_dist_knn.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int, C.c_float, C.c_int, C.c_int,
                                     C.c_void_p, C.c_void_p, C.c_void_p]
    sim.b200knn_dist_knn.restype = C.c_int
    x = gen.uniform(300, seed=2)
    out = np.zeros(300, np.float32)
    assert sim.b200knn_dist_knn(eng, x.ctypes.data, 300, 0, 0, math.inf, 2, 0, out.ctypes.data, None, None) != 0
    assert "k must be >= 1" in err(sim)
    assert sim.b200knn_dist_knn(eng, x.ctypes.data, 300, 0, 4, math.inf, 9, 0, out.ctypes.data, None, None) != 0
    assert "unknown mode" in err(sim)
    # global point ids must fit 32 bits in redistribute mode: a slice that starts beyond 2^32 is refused
    assert sim.b200knn_dist_knn(eng, x.ctypes.data, 300, 2**32, 4, math.inf, 2, 0, out.ctypes.data, None, None) != 0
    assert "global point ids must fit 32 bits" in err(sim)
    assert sim.b200knn_dist_knn(eng, x.ctypes.data, 300, 0, 4, math.inf, 2, 0, out.ctypes.data, None, None) == 0
    sim.b200knn_dist_destroy.argtypes = [C.c_void_p]
    sim.b200knn_dist_destroy(eng)


def test_launcher_reports_a_failing_rank(tmp_path):
    """knnrun exits with the first non-zero exit code of its ranks (one rank's missing per-rank file here) and
    does not hang the others in a collective."""
    run, pv = str(ROOT / "build" / "knnrun"), str(ROOT / "build" / "cudaMpiKNN_prePartitionedData")
    a = tmp_path / "a.float3"
    kio.write_float3(str(a), gen.uniform(200, seed=3))
    lst = tmp_path / "files.txt"
    lst.write_text(f"{a}\n{a}\n{a}\n")                       # 3 files for 2 ranks: the program must refuse
    env = dict(os.environ, B200KNN_BACKEND="cpu")
    r = subprocess.run([run, "-n", "2", pv, str(lst), "-o", str(tmp_path / "out"), "-k", "4"], env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=120)
    assert r.returncode != 0
    assert "number of ranks" in (r.stderr + r.stdout) or "files" in (r.stderr + r.stdout)
