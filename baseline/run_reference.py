"""bench.py --impl reference: run the UNMODIFIED reference (baseline/_ref) for the same metric."""
import json
import os
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]


def bench_reference(args) -> int:
    if int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps({"impl": "reference", "unavailable": "reference build (MPI shim) not wired up yet"}), flush=True)
    return 0
