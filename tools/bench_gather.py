#!/usr/bin/env python
"""Times the final coefficient gather of BASELINE configs[3] (fp64 (6, 4096, 4096) map, 805 MB, rows split
over the ranks) over NCCL in both forms of sharding.gather_planar: point-to-point group and all-gather + permute.

    torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/bench_gather.py
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from rti_b200 import sharding  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
    R = 4096
    r0, nr = sharding.row_partition(R, world)[rank]
    local = torch.full((6, nr * R), float(rank), dtype=torch.float64, device="cuda")
    out = {}
    for method in ("p2p", "allgather"):
        for _ in range(2):
            full = sharding.gather_planar(local, R, R, dst=0, method=method)
        dist.barrier()
        torch.cuda.synchronize()
        ts = []
        for _ in range(5):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            dist.barrier()
            a.record()
            full = sharding.gather_planar(local, R, R, dst=0, method=method)
            b.record()
            torch.cuda.synchronize()
            t = torch.tensor([a.elapsed_time(b)], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ts.append(float(t.item()))
        if rank == 0:
            ok = all(bool((full[:, s0:s0 + sn] == float(r)).all()) for r, (s0, sn) in enumerate(sharding.row_partition(R, world)))
            nbytes = 48 * R * R * (world - 1) // world
            out[method] = {"ms_best": min(ts), "ms_all": ts, "bytes_into_rank0": nbytes,
                           "gbs": nbytes / (min(ts) * 1e-3) / 1e9, "correct": ok}
        del full
    if rank == 0:
        print(json.dumps({"ranks": world, "gather": out}))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
