"""Rank plumbing for the z-slab decomposition (one process per GPU).

The halo exchange itself happens inside libfdtd3d_b200.so (ncclSend / ncclRecv on the solver's communication
stream).  What has to happen outside is only the rendezvous: rank 0 creates the 128-byte NCCL unique id and every
other rank must receive it before fdtd_comm_init.  This module does that over torch.distributed (any backend:
nccl on the GPU box, gloo in the CPU tests) and offers the small reductions bench.py needs.
"""
import os

import numpy as np

from . import Solver, slab_range  # noqa: F401


def env_rank():
    """(rank, world_size, local_rank) from the torchrun environment; (0, 1, 0) when not launched by it."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0")))


def share_unique_id(dist, rank, device, make_id=None):
    """Rank 0 obtains the 128-byte id from `make_id` (default: fdtd_comm_unique_id) and broadcasts it."""
    import torch
    buf = torch.zeros(128, dtype=torch.uint8, device=device)
    if rank == 0:
        raw = (make_id or Solver.comm_unique_id)()
        assert len(raw) == 128
        buf.copy_(torch.frombuffer(bytearray(raw), dtype=torch.uint8))
    dist.broadcast(buf, 0)
    return bytes(buf.cpu().numpy().tobytes())


def max_over_ranks(dist, value, device):
    """Max of a python float over all ranks (multi-GPU timings are the max over ranks)."""
    import torch
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def slab_of_array(array, shape, z0, nz):
    """The planes [z0, z0+nz) of a reference-order (i, j, k) array as a contiguous reference-order block."""
    return np.ascontiguousarray(np.asarray(array).reshape(shape)[:, :, z0:z0 + nz]).reshape(-1)
