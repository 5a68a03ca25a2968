"""Parity with the reference's own binaries (built unmodified from /root/reference by
baseline/build_ref.sh, run through baseline/mpi_shim): identical OUTPUT FILES, byte for byte.
This is BASELINE.json config 1 (unordered, 1 M uniform, k = 10, 2 ranks) plus a pre-partitioned case."""
import math
import os
import sys
from pathlib import Path

import numpy as np
import pytest

torch = pytest.importorskip("torch")

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

from baseline import run_reference as ref  # noqa: E402
from mpi_cuda_largescaleknn_b200.utils import generators as gen  # noqa: E402
from mpi_cuda_largescaleknn_b200.utils import io as kio  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def built():
    err = ref.ensure_built()
    if err:
        pytest.skip(err)
    return True


def _ours(x, k, r=math.inf):
    from mpi_cuda_largescaleknn_b200.ops.knn import knn
    return knn(torch.from_numpy(x).to("cuda:0"), k, r).cpu().numpy()


def test_unordered_output_file_identical_to_reference(built, tmp_path):
    n, k = 1_000_000, 10
    x = gen.uniform(n, seed=42)
    inp, out = tmp_path / "in.float3", tmp_path / "out.float"
    kio.write_float3(str(inp), x)
    res, why = ref.run_once(ref.EXE_UNORDERED, str(inp), str(out), k, ranks=2, gpus=1, timeout_s=600)
    assert res is not None, why
    theirs = kio.read_floats(str(out))
    assert theirs.shape == (n,)
    mine = _ours(x, k)
    assert np.array_equal(theirs.view(np.uint32), mine.view(np.uint32))


def test_prepartitioned_output_files_identical_to_reference(built, tmp_path):
    parts = gen.gaussian_blobs(60_000, 2, overlap=0.10, seed=3)
    names = []
    for i, p in enumerate(parts):
        f = tmp_path / f"part{i}.float3"
        kio.write_float3(str(f), p)
        names.append(str(f))
    lst = tmp_path / "files.txt"
    lst.write_text("\n".join(names) + "\n")
    prefix = tmp_path / "dist"
    res, why = ref.run_once(ref.EXE_PREPART, str(lst), str(prefix), 16, ranks=2, gpus=1, timeout_s=600)
    assert res is not None, why
    x_all = np.concatenate(parts)
    mine = _ours(x_all, 16)
    off = 0
    for i, p in enumerate(parts):
        theirs = kio.read_floats(kio.per_rank_output_name(str(prefix), i))
        assert np.array_equal(theirs.view(np.uint32), mine[off:off + len(p)].view(np.uint32))
        off += len(p)
