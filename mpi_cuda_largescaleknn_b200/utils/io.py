"""Binary file formats of the reference, unchanged.

* input: raw little-endian float32 triples, no header, ``count = file_size // 12``
  (/root/reference/unorderedDataVariant.cu:41-63).
* unordered slice of rank r of R: ``[count*r//R, count*(r+1)//R)`` (same file, :56-59).
* pre-partitioned: a text file with one input path per line (prePartitionedDataVariant.cu:114-126).
  Divergence, on purpose: a last line without trailing newline IS accepted here (the reference
  drops it silently) and blank lines are ignored.
* output: float32 per input point; per-rank files are ``<prefix>_%06d.float``
  (prePartitionedDataVariant.cu:380-385).
"""
from __future__ import annotations

import os
from typing import List, Tuple

import numpy as np


def count_points(path: str) -> int:
    try:
        return os.path.getsize(path) // 12
    except OSError:
        return 0   # the reference treats a missing file as empty


def slice_bounds(count: int, rank: int, size: int) -> Tuple[int, int]:
    return count * rank // size, count * (rank + 1) // size


def read_float3(path: str, rank: int = 0, size: int = 1) -> Tuple[np.ndarray, int, int]:
    """Return (points[n,3] float32, begin, total_count) for this rank's slice."""
    total = count_points(path)
    begin, end = slice_bounds(total, rank, size)
    if end == begin:
        return np.zeros((0, 3), np.float32), begin, total
    with open(path, "rb") as f:
        f.seek(begin * 12)
        buf = np.fromfile(f, dtype="<f4", count=(end - begin) * 3)
    return buf.reshape(-1, 3), begin, total


def write_float3(path: str, pts: np.ndarray) -> None:
    np.ascontiguousarray(pts, dtype="<f4").tofile(path)


def read_file_list(path: str) -> List[str]:
    with open(path, "r") as f:
        return [ln.strip() for ln in f.read().split("\n") if ln.strip()]


def per_rank_output_name(prefix: str, rank: int) -> str:
    return "%s_%06d.float" % (prefix, rank)


def write_floats_at(path: str, values: np.ndarray, begin: int, total: int, create: bool) -> None:
    """Write this rank's slice of the single output file at its offset (no barrier chain needed;
    produces the same bytes as the reference's rank-ordered append, unorderedDataVariant.cu:229-237)."""
    if create:
        with open(path, "wb") as f:
            f.truncate(total * 4)
    with open(path, "r+b") as f:
        f.seek(begin * 4)
        np.ascontiguousarray(values, dtype="<f4").tofile(f)


def read_floats(path: str) -> np.ndarray:
    return np.fromfile(path, dtype="<f4")
