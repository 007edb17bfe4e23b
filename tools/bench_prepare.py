#!/usr/bin/env python
"""prepare_images_data (analysis.py:375-411) alone through the drop-in: a PTM result of a rows x cols ROI on
the device -> the host table the reference returns (list[100] of list[100] of int32 (R, R) arrays), with the
tiles leaving through rti_table_tile_to_host ("native") and through the earlier Python route ("python").

    python tools/bench_prepare.py [--size 1024] [--n 64]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import rti_b200  # noqa: E402
from rti_b200 import synthetic  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=1024)
    ap.add_argument("--n", type=int, default=64)
    args = ap.parse_args()
    torch.cuda.set_device(0)
    r = args.size
    rti_b200.constants.ROI_DIAMETER = r
    frames, cams = synthetic.random_stack(r, r, args.n, seed=0)
    data = rti_b200.compute_intensities(synthetic.as_records(frames, cams))
    res = rti_b200.interpolate_intensities(data, interpolate_PTM=True)
    small = None
    for route in ("native", "python", "native"):
        os.environ["RTI_TABLE_TO_HOST"] = route
        t0 = time.perf_counter()
        table = rti_b200.prepare_images_data(res)
        dt = time.perf_counter() - t0
        chk = [int(table[v][u][::7, ::5].astype(np.int64).sum()) for v, u in ((0, 0), (3, 7), (50, 50), (99, 99))]
        small = chk if small is None else small
        print(json.dumps({"route": route, "s": dt, "table_gb": 4e-9 * 10000 * r * r, "gbs": 4e-9 * 10000 * r * r / dt,
                          "same_values": chk == small}), flush=True)
        del table


if __name__ == "__main__":
    main()
