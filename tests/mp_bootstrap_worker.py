"""CPU worker (gloo, world_size > 1): the torch.distributed bootstrap that the C++ driver calls back
into — allgather of equal-sized host blobs and barrier — plus the reference's index split."""
import ctypes as C
import os
import sys
from pathlib import Path

import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))

from mpi_cuda_largescaleknn_b200.parallel.dist import TorchBootstrap, global_slice  # noqa: E402
from mpi_cuda_largescaleknn_b200.parallel.ring import allgather_bytes, allreduce_max_int  # noqa: E402


def main() -> int:
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    bs = TorchBootstrap()
    assert (bs.rank, bs.world) == (rank, world)
    # through the C callback signatures, exactly as libb200knn calls them
    for nbytes in (1, 24, 256, 4096):
        send = (C.c_ubyte * nbytes)(*([(rank * 7 + i) % 251 for i in range(nbytes)]))
        recv = (C.c_ubyte * (nbytes * world))()
        rc = bs.struct.allgather(None, C.cast(send, C.c_void_p), C.cast(recv, C.c_void_p), nbytes)
        assert rc == 0
        for r in range(world):
            assert list(recv[r * nbytes:(r + 1) * nbytes]) == [(r * 7 + i) % 251 for i in range(nbytes)]
    assert bs.struct.barrier(None) == 0
    blobs = allgather_bytes(bytes([rank]) * 16)
    assert [b[0] for b in blobs] == list(range(world))
    assert allreduce_max_int(100 + rank) == 100 + world - 1
    # slices tile the global range exactly like the reference's readFilePortion
    total = 1000003
    b, e = global_slice(total, rank, world)
    ends = allgather_bytes(e.to_bytes(8, "little"))
    if rank > 0:
        assert int.from_bytes(ends[rank - 1], "little") == b
    if rank == world - 1:
        assert e == total
    dist.barrier()
    if rank == 0:
        print("MP_BOOTSTRAP_OK", flush=True)
    dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
