"""world_size-2 CPU (gloo) test of the N>1 plumbing: row partition + final gather."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, rows, cols, out_path):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    from rti_b200 import sharding
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        r0, nr = sharding.row_partition(rows, world)[rank]
        # a "coefficient map" whose value encodes (plane, global row, col)
        k = torch.arange(6, dtype=torch.float64).view(6, 1, 1) * 1e6
        y = torch.arange(r0, r0 + nr, dtype=torch.float64).view(1, nr, 1) * 1e3
        x = torch.arange(cols, dtype=torch.float64).view(1, 1, cols)
        local = (k + y + x).reshape(6, nr * cols)
        full = sharding.gather_planar(local, rows, cols, dst=0)
        if rows % world == 0 and rows > 0:
            # the all-gather form (equal slabs) gives the same map
            full2 = sharding.gather_planar(local, rows, cols, dst=0, method="allgather")
            assert (full2 is None) == (rank != 0)
            if rank == 0:
                assert torch.equal(full, full2)
        if rank == 0:
            np.save(out_path, full.numpy())
        else:
            assert full is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("rows,world", [(7, 2), (8, 2), (1, 2)])
def test_gather_planar_gloo(tmp_path, rows, world):
    cols = 5
    out = str(tmp_path / "full.npy")
    mp.spawn(_worker, args=(world, _free_port(), rows, cols, out), nprocs=world, join=True)
    full = np.load(out)
    k = np.arange(6).reshape(6, 1, 1) * 1e6
    y = np.arange(rows).reshape(1, rows, 1) * 1e3
    x = np.arange(cols).reshape(1, 1, cols)
    assert np.array_equal(full, k + y + x)
