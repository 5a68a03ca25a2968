"""CPU-only: file formats, generators and pretty printers."""
import numpy as np

from mpi_cuda_largescaleknn_b200.utils import generators as gen
from mpi_cuda_largescaleknn_b200.utils import io as kio
from mpi_cuda_largescaleknn_b200.utils.pretty import pretty_bytes, pretty_number, pretty_seconds


def test_float3_roundtrip_and_slices(tmp_path):
    x = gen.uniform(1003, seed=1)
    p = tmp_path / "a.float3"
    kio.write_float3(str(p), x)
    assert kio.count_points(str(p)) == 1003
    parts, begins = [], []
    for r in range(7):
        pts, b, total = kio.read_float3(str(p), r, 7)
        assert total == 1003 and b == 1003 * r // 7
        parts.append(pts)
        begins.append(b)
    assert np.array_equal(np.concatenate(parts), x)
    assert kio.count_points(str(tmp_path / "missing.float3")) == 0


def test_single_output_file_written_by_slices(tmp_path):
    vals = np.arange(100, dtype=np.float32)
    out = tmp_path / "o.float"
    kio.write_floats_at(str(out), vals[:40], 0, 100, create=True)
    kio.write_floats_at(str(out), vals[70:], 70, 100, create=False)
    kio.write_floats_at(str(out), vals[40:70], 40, 100, create=False)
    assert np.array_equal(kio.read_floats(str(out)), vals)
    assert kio.per_rank_output_name("pre", 7) == "pre_000007.float"


def test_file_list_accepts_missing_trailing_newline(tmp_path):
    lst = tmp_path / "l.txt"
    lst.write_text("a.float3\n\nb.float3")
    assert kio.read_file_list(str(lst)) == ["a.float3", "b.float3"]


def test_generators_shapes_and_determinism():
    assert gen.uniform(10, seed=3).shape == (10, 3)
    assert np.array_equal(gen.uniform(10, seed=3), gen.uniform(10, seed=3))
    blobs = gen.gaussian_blobs(100, 8, overlap=0.1, seed=1)
    assert len(blobs) == 8 and all(b.shape == (100, 3) and b.dtype == np.float32 for b in blobs)
    c = gen.clustered_surface(1001, seed=2)
    assert c.shape == (1001, 3) and c.dtype == np.float32


def test_pretty_printers():
    assert pretty_number(1234567) == "1.23M"
    assert pretty_number(999) == "999"
    assert pretty_bytes(1536) == "1.50KiB"
    assert pretty_seconds(0.0123) == "12.30ms"
