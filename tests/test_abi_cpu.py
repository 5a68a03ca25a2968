"""CPU-only: the three descriptions of the C ABI agree — the header, the shared library's exports,
and the ctypes mirror structures the Python package passes across the boundary."""
import ctypes as C
import re
import subprocess
import sys
from pathlib import Path

import pytest

from conftest import run_make

ROOT = Path(__file__).resolve().parents[1]
HEADER = ROOT / "include" / "b200knn.h"


def _declared():
    text = re.sub(r"/\*.*?\*/", "", HEADER.read_text(), flags=re.S)
    return set(re.findall(r"\b(b200knn_[a-z0-9_]+)\s*\(", text))


def _exported():
    from mpi_cuda_largescaleknn_b200.ops._lib import LIB_PATH, lib
    lib()
    out = subprocess.run(["nm", "-D", "--defined-only", str(LIB_PATH)], check=True, stdout=subprocess.PIPE, text=True).stdout
    return {line.split()[-1] for line in out.splitlines() if " T " in line and "b200knn_" in line and "_Z" not in line}


def test_every_declared_function_is_exported_and_vice_versa():
    declared, exported = _declared(), _exported()
    assert declared - exported == set(), f"declared in the header but not in the library: {sorted(declared - exported)}"
    assert exported - declared == set(), f"exported but missing from the header: {sorted(exported - declared)}"


def test_python_only_calls_functions_the_header_declares():
    declared = _declared()
    used = set()
    for py in (ROOT / "mpi_cuda_largescaleknn_b200").rglob("*.py"):
        used |= set(re.findall(r"\b(b200knn_[a-z0-9_]+)\b", py.read_text()))
    used = {u for u in used if u not in ("b200knn_step_args", "b200knn_dist_report", "b200knn_bootstrap")}
    assert used - declared == set(), sorted(used - declared)


def test_ctypes_structures_have_the_layout_of_the_c_structs(tmp_path):
    """Compile a few sizeof/offsetof probes with the host compiler and compare with ctypes."""
    from mpi_cuda_largescaleknn_b200.ops._lib import StepArgs
    from mpi_cuda_largescaleknn_b200.parallel.dist import Bootstrap, DistReport
    probes = {
        "b200knn_step_args": (StepArgs, ["max_radius", "inbox_heap", "outbox_cnt", "result_sorted"]),
        "b200knn_dist_report": (DistReport, ["pending_groups", "chunks", "ms_local_phase", "auto_chose_redistribute"]),
        "b200knn_bootstrap": (Bootstrap, ["world", "allgather", "barrier"]),
    }
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "b200knn.h"', "int main(void) {"]
    for cname, (_, fields) in probes.items():
        lines.append(f'  printf("{cname} %zu\\n", sizeof({cname}));')
        for f in fields:
            lines.append(f'  printf("{cname}.{f} %zu\\n", offsetof({cname}, {f}));')
    lines += ["  return 0;", "}"]
    src = tmp_path / "probe.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "probe"
    cc = subprocess.run(["gcc", "-std=c99", "-I", str(ROOT / "include"), str(src), "-o", str(exe)],
                        stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert cc.returncode == 0, cc.stdout          # also proves the header is plain C
    got = dict(line.split() for line in subprocess.run([str(exe)], check=True, stdout=subprocess.PIPE, text=True).stdout.splitlines())
    for cname, (ctype, fields) in probes.items():
        assert int(got[cname]) == C.sizeof(ctype), cname
        for f in fields:
            assert int(got[f"{cname}.{f}"]) == getattr(ctype, f).offset, f"{cname}.{f}"


def test_header_constants_match_python():
    from mpi_cuda_largescaleknn_b200.ops import _lib
    from mpi_cuda_largescaleknn_b200.parallel.dist import MODES
    text = HEADER.read_text()
    assert int(re.search(r"#define\s+B200KNN_MAX_RANKS\s+(\d+)", text).group(1)) == _lib.MAX_RANKS
    assert int(re.search(r"#define\s+B200KNN_IPC_BLOB_BYTES\s+(\d+)", text).group(1)) == _lib.IPC_BLOB_BYTES
    for name, val in MODES.items():
        m = re.search(rf"B200KNN_MODE_{name.upper()}\s*=\s*(\d+)", text)
        assert m and int(m.group(1)) == val, name



def test_c_example_builds_against_the_library():
    """examples/sample_knn.c is plain C99 against include/b200knn.h + libb200knn.so (run it on a GPU box)."""
    out = run_make("examples")
    assert out.returncode == 0, out.stdout[-2000:]
    assert (ROOT / "build" / "sample_knn").exists()


if __name__ == "__main__":
    sys.exit(pytest.main([__file__, "-q"]))
