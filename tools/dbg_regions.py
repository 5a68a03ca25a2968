import os, sys, numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.getcwd())
from mpi_cuda_largescaleknn_b200.parallel.dist import DistKnn
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank); dev = torch.device("cuda", rank)
dist.init_process_group("cpu:gloo,cuda:nccl", device_id=dev)
eng = DistKnn()
n = 670000
g = torch.Generator(device=dev); g.manual_seed(100 + rank)
sigma = 0.5 / 2.128
x = torch.randn((n, 3), generator=g, device=dev) * sigma + torch.tensor([0.0, 0.0, float(rank)], device=dev)
for mode in ("redistribute", "prepartitioned"):
    out = eng.knn(x, 100, mode=mode, global_begin=rank * n)
    rep = eng.last_report
    t = torch.tensor([rep["pending_groups"], rep["points_after_redistribute"]], device=dev, dtype=torch.float64)
    dist.all_reduce(t)
    if rank == 0:
        print(mode, "pending groups total", int(t[0].item()), "of", 2 * n // 32, "points after", rep["points_after_redistribute"], flush=True)
eng.close(); dist.barrier(); dist.destroy_process_group()
