#!/usr/bin/env python
"""Smallest case that touches every single-GPU kernel: for compute-sanitizer (racecheck / memcheck)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import ctypes as C
import numpy as np
import torch
from mpi_cuda_largescaleknn_b200.ops._lib import lib
from mpi_cuda_largescaleknn_b200.ops.knn import KnnIndex, fcp
from mpi_cuda_largescaleknn_b200.utils.oracle import bit_equal, bruteforce

n = int(sys.argv[1]) if len(sys.argv) > 1 else 6000
x = np.random.default_rng(0).random((n, 3), dtype=np.float32)
xd = torch.from_numpy(x).cuda()
idx = KnnIndex(xd)
a = idx.knn_self(10).cpu().numpy()
q = np.random.default_rng(1).random((777, 3), dtype=np.float32)
b = idx.knn_query(torch.from_numpy(q).cuda(), 5).cpu().numpy()
ids, d = fcp(idx, torch.from_numpy(q).cuda())
out = torch.empty(n, device="cuda")
L = lib()
L.b200knn_knn_self_noalloc.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_void_p, C.c_void_p]
assert L.b200knn_knn_self_noalloc(xd.data_ptr(), n, 10, float("inf"), out.data_ptr(), None) == 0
torch.cuda.synchronize()
ok = bit_equal(a, bruteforce(x, k=10)) and bit_equal(b, bruteforce(x, queries=q, k=5)) and \
    bit_equal(d.cpu().numpy(), bruteforce(x, queries=q, k=1)) and bit_equal(out.cpu().numpy(), a)
idx.close()
print("SANITIZE_CASE_OK" if ok else "SANITIZE_CASE_MISMATCH")
sys.exit(0 if ok else 1)
