"""Host-side logic of the multi-GPU path with world_size 2 on CPU (gloo): the unique-id rendezvous, the slab
partition each rank derives, the slab extraction used to compare against a whole-grid result, and max-over-ranks."""
import os
import socket
import sys

import numpy as np
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    import fdtd3d_b200 as F
    from fdtd3d_b200 import multi
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    fake = bytes(range(128))
    got = multi.share_unique_id(dist, rank, "cpu", make_id=lambda: fake)
    assert got == fake
    nz = 27
    z0, nzl = F.slab_range(nz, rank, world)
    full = np.arange(4 * 3 * nz, dtype=np.float64)
    slab = multi.slab_of_array(full, (4, 3, nz), z0, nzl)
    assert slab.shape == (4 * 3 * nzl,)
    assert np.array_equal(slab.reshape(4, 3, nzl), full.reshape(4, 3, nz)[:, :, z0:z0 + nzl])
    m = multi.max_over_ranks(dist, 10.0 + rank, "cpu")
    assert m == 10.0 + world - 1
    # every rank reports its slab; rank 0 checks that the slabs tile [0, nz) in order
    import torch
    t = torch.tensor([z0, nzl], dtype=torch.int64)
    parts = [torch.zeros(2, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(parts, t)
    if rank == 0:
        pos = 0
        for p in parts:
            assert int(p[0]) == pos
            pos += int(p[1])
        assert pos == nz
        open(os.path.join(out_dir, "ok"), "w").write("ok")
    dist.destroy_process_group()


def test_rendezvous_and_partition_world2(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok").exists()
