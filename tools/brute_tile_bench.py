#!/usr/bin/env python
"""SURVEY.md section 8 row 10, measured: the CUDA-core leaf sweep against a TMA + tcgen05 (tf32, TMEM) conservative
filter on the same inputs (csrc/brute_tile.cu).

    python tools/brute_tile_bench.py [--nq 1048576] [--np 2048] [--k 100]

Prints one JSON line: time of either kernel (CUDA events, after warm-up), pairs per second, whether the tensor-core
masks are supersets of the exact ones on a sample, and how many more candidates the filter lets through."""
import argparse, ctypes as C, json, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from mpi_cuda_largescaleknn_b200.ops._lib import lib, check


def run(nq, np_, k, reps=10, sample=4096, seed=0, scale=0.05):
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev); g.manual_seed(seed)
    centre = torch.tensor([0.37, 0.61, 0.52], device=dev)
    p = centre + (torch.rand((np_, 3), generator=g, device=dev) - 0.5) * scale       # one neighbourhood of the cloud
    q = centre + (torch.rand((nq, 3), generator=g, device=dev) - 0.5) * scale
    thr = torch.empty(nq, device=dev)
    for b in range(0, nq, 65536):                                                     # k-th smallest squared distance
        d2 = torch.cdist(q[b:b + 65536], p) ** 2
        thr[b:b + 65536] = d2.kthvalue(min(k, np_), dim=1).values
    L = lib()
    vp, i64, i32, f32 = C.c_void_p, C.c_int64, C.c_int, C.c_float
    L.b200knn_brute_tile_prep.argtypes = [vp, i64, C.POINTER(C.c_float * 3), i32, vp, vp, vp]
    L.b200knn_brute_tile_cuda.argtypes = [vp, i64, vp, i32, vp, vp, vp, vp]
    L.b200knn_brute_tile_tc.argtypes = [vp, vp, i64, vp, i32, f32, vp, vp, vp, i32, vp]
    for fn in (L.b200knn_brute_tile_prep, L.b200knn_brute_tile_cuda, L.b200knn_brute_tile_tc):
        fn.restype = i32
    st = torch.cuda.current_stream().cuda_stream
    c3 = (C.c_float * 3)(*[float(v) for v in centre.cpu()])
    q8, p8 = torch.empty((nq, 8), device=dev), torch.empty((np_, 8), device=dev)
    qn, pn = torch.empty(nq, device=dev), torch.empty(np_, device=dev)
    ev = lambda: torch.cuda.Event(enable_timing=True)
    e0, e1 = ev(), ev()
    e0.record()
    check(L.b200knn_brute_tile_prep(q.data_ptr(), nq, C.byref(c3), 1, q8.data_ptr(), qn.data_ptr(), st), "prep q")
    check(L.b200knn_brute_tile_prep(p.data_ptr(), np_, C.byref(c3), 0, p8.data_ptr(), pn.data_ptr(), st), "prep p")
    e1.record(); torch.cuda.synchronize()
    prep_ms = e0.elapsed_time(e1)
    pmax = float(pn.max().item())
    cnt_c = torch.zeros(nq, dtype=torch.int32, device=dev)
    cnt_t = torch.zeros(nq, dtype=torch.int32, device=dev)
    sms = torch.cuda.get_device_properties(0).multi_processor_count

    def cuda_once(masks=None):
        check(L.b200knn_brute_tile_cuda(q.data_ptr(), nq, p.data_ptr(), np_, thr.data_ptr(), cnt_c.data_ptr(),
                                        masks.data_ptr() if masks is not None else None, st), "cuda")

    def tc_once(masks=None):
        check(L.b200knn_brute_tile_tc(q8.data_ptr(), qn.data_ptr(), nq, p8.data_ptr(), np_, pmax, thr.data_ptr(),
                                      cnt_t.data_ptr(), masks.data_ptr() if masks is not None else None, 4 * sms, st), "tc")

    def timed(fn):
        fn(); torch.cuda.synchronize()
        a, b = ev(), ev()
        a.record()
        for _ in range(reps):
            fn()
        b.record(); torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    ms_c, ms_t = timed(cuda_once), timed(tc_once)
    # masks on a sample: superset check
    ns = min(sample, nq)
    words = (np_ + 31) // 32
    mc = torch.zeros((nq, words), dtype=torch.int32, device=dev)
    mt = torch.zeros((nq, words), dtype=torch.int32, device=dev)
    cuda_once(mc); tc_once(mt); torch.cuda.synchronize()
    dropped = int(((mc[:ns] & ~mt[:ns]) != 0).sum().item())
    exact_want = (torch.cdist(q[:ns], p) ** 2 < thr[:ns, None]).sum(1)
    pairs = nq * np_
    return {"nq": nq, "np": np_, "k": k, "cuda_core_ms": ms_c, "tcgen05_ms": ms_t, "prep_ms_once": prep_ms,
            "cuda_core_Gpairs_per_s": pairs / ms_c / 1e6, "tcgen05_Gpairs_per_s": pairs / ms_t / 1e6,
            "speedup_filter_only": ms_c / ms_t, "true_candidates_dropped_by_filter": dropped,
            "candidates_exact": int(cnt_c.sum().item()), "candidates_filter": int(cnt_t.sum().item()),
            "filter_inflation": float(cnt_t.sum().item()) / max(1, int(cnt_c.sum().item())),
            "cuda_counts_vs_torch_max_abs_diff": int((cnt_c[:ns].long() - exact_want).abs().max().item())}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--nq", type=int, default=1 << 20)
    ap.add_argument("--np", dest="np_", type=int, default=2048)
    ap.add_argument("--k", type=int, default=100)
    a = ap.parse_args()
    print(json.dumps(run(a.nq, a.np_, a.k)))
