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


def test_host_prefetcher_schedule_with_fake_streams(monkeypatch):
    """Control flow of ops.pipeline.HostPrefetcher without a GPU: streams and events are recorded as a trace.
    Cloud i must be yielded from slot i % 2, its upload must wait for the computation that last used the slot,
    and the upload of cloud i+1 must be issued before cloud i is handed to the consumer."""
    import types

    import torch

    import mpi_cuda_largescaleknn_b200.ops.pipeline as pl
    trace = []

    class Event:
        def __init__(self):
            self.name = None

        def record(self, stream=None):
            trace.append(("record", self.name))

    class Stream:
        def __init__(self, device=None, name="copy"):
            self.name = name

        def wait_event(self, e):
            trace.append((self.name + " waits", e.name))

    class Ctx:
        def __init__(self, s):
            pass

        def __enter__(self):
            return self

        def __exit__(self, *a):
            return False

    fake_cuda = types.SimpleNamespace(Stream=Stream, Event=Event, stream=Ctx, current_stream=lambda d=None: Stream(name="main"))
    monkeypatch.setattr(pl, "torch", types.SimpleNamespace(cuda=fake_cuda, empty=torch.empty, float32=torch.float32,
                                                         device=torch.device, Tensor=torch.Tensor))
    p = pl.HostPrefetcher(4, "cpu")
    for i in range(2):
        p.ready[i].name, p.free[i].name = f"ready{i}", f"free{i}"
    clouds = [torch.full((4, 3), float(i)) for i in range(4)]
    for c in clouds:
        c.is_pinned = lambda: True
    seen = []
    for i, staged in enumerate(p.feed(clouds)):
        trace.append(("compute", i))
        seen.append(float(staged[0, 0]))
    assert seen == [0.0, 1.0, 2.0, 3.0]
    # the upload of cloud 2 (slot 0 again) waits for the computation on cloud 0, and is issued before compute 1
    assert trace.index(("copy waits", "free0")) < trace.index(("compute", 0))          # first use: nothing to wait for yet
    second_wait = [j for j, t in enumerate(trace) if t == ("copy waits", "free0")][1]
    assert trace.index(("record", "free0")) < second_wait < trace.index(("compute", 1))


def test_result_drain_schedule_with_fake_streams(monkeypatch):
    """Control flow of ops.pipeline.ResultDrain without a GPU: step i writes slot i % 2, its download waits for the
    step, the next step that writes the same slot waits for that download, and submit() hands back the result of the
    step before, after waiting for ITS download only."""
    import types

    import torch

    import mpi_cuda_largescaleknn_b200.ops.pipeline as pl
    trace = []

    class Event:
        def __init__(self):
            self.name = None

        def record(self, stream=None):
            trace.append(("record", self.name))

        def synchronize(self):
            trace.append(("host waits", self.name))

    class Stream:
        def __init__(self, device=None, name="copy"):
            self.name = name

        def wait_event(self, e):
            trace.append((self.name + " waits", e.name))

    class Ctx:
        def __init__(self, s):
            pass

        def __enter__(self):
            return self

        def __exit__(self, *a):
            return False

    fake_cuda = types.SimpleNamespace(Stream=Stream, Event=Event, stream=Ctx, current_stream=lambda d=None: Stream(name="main"))
    real_empty = torch.empty
    monkeypatch.setattr(pl, "torch", types.SimpleNamespace(
        cuda=fake_cuda, float32=torch.float32, device=torch.device, Tensor=torch.Tensor,
        empty=lambda *a, pin_memory=False, **kw: real_empty(*a, **kw)))
    d = pl.ResultDrain(4, "cpu")
    for i in range(2):
        d.filled[i].name, d.done[i].name = f"filled{i}", f"done{i}"
    got = []
    for i in range(4):
        out = d.next_out()
        out.fill_(float(i))                    # "the step"
        trace.append(("compute", i))
        prev = d.submit()
        got.append(None if prev is None else float(prev[0]))
    last = d.finish()
    assert got == [None, 0.0, 1.0, 2.0] and float(last[0]) == 3.0
    # step 2 reuses slot 0: the main stream waits for download 0 before computing, and the host waited only for done0
    c2 = trace.index(("compute", 2))
    waits0 = [j for j, t in enumerate(trace) if t == ("main waits", "done0")]
    assert len(waits0) == 2 and trace.index(("record", "done0")) < waits0[1] < c2
    assert trace.index(("copy waits", "filled0")) > trace.index(("compute", 0))
    # handing in step 1 returns step 0 after waiting for done0 — never for done1, which is still in flight
    s1 = trace.index(("compute", 1))
    assert ("host waits", "done0") in trace[s1:c2] and ("host waits", "done1") not in trace[s1:c2]
