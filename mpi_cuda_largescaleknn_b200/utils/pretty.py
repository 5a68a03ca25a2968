"""Human-readable numbers and a wall-clock helper.

Role of ``prettyDouble`` / ``prettyNumber`` / ``prettyBytes`` / ``getCurrentTime`` in the reference's
library (cudaKDTree/cukd/common.h:157-237)."""
from __future__ import annotations

import time


def pretty_number(n: float) -> str:
    """1234567 -> '1.23M' (decimal prefixes, as cukd's prettyNumber)."""
    n = float(n)
    for div, suffix in ((1e12, "T"), (1e9, "G"), (1e6, "M"), (1e3, "K")):
        if abs(n) >= div:
            return f"{n / div:.2f}{suffix}"
    return f"{n:.0f}" if n == int(n) else f"{n:.2f}"


def pretty_bytes(n: float) -> str:
    """Binary prefixes: 1536 -> '1.50KiB'."""
    n = float(n)
    for div, suffix in ((1 << 40, "TiB"), (1 << 30, "GiB"), (1 << 20, "MiB"), (1 << 10, "KiB")):
        if abs(n) >= div:
            return f"{n / div:.2f}{suffix}"
    return f"{n:.0f}B"


def pretty_seconds(s: float) -> str:
    s = float(s)
    if s >= 1.0:
        return f"{s:.2f}s"
    if s >= 1e-3:
        return f"{s * 1e3:.2f}ms"
    if s >= 1e-6:
        return f"{s * 1e6:.2f}us"
    return f"{s * 1e9:.0f}ns"


def current_time() -> float:
    """Seconds, monotonic (cukd's getCurrentTime uses gettimeofday)."""
    return time.perf_counter()
