#!/usr/bin/env python
"""Minimal driver for ncu captures of the fused PTM kernel (warm-up + timed launches).

    python tools/prof_fused.py [--rows 512] [--cols 1024] [--n 500] [--reps 3]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from rti_b200 import _capi, synthetic  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=512)
    ap.add_argument("--cols", type=int, default=1024)
    ap.add_argument("--n", type=int, default=500)
    ap.add_argument("--reps", type=int, default=3)
    args = ap.parse_args()
    torch.cuda.set_device(0)
    _capi.load()
    rows, cols, n, g = args.rows, args.cols, args.n, 100
    frames, cams = synthetic.random_stack(rows, cols, n, seed=0)
    d_f, d_c = torch.from_numpy(frames).cuda(), torch.from_numpy(cams).cuda()
    p = rows * cols
    axis = torch.from_numpy(np.around(np.mgrid[-1:1:0.02], 2)).cuda()
    coef = torch.empty((6, p), dtype=torch.float64, device="cuda")
    table = torch.empty((g * g, p), dtype=torch.int32, device="cuda")
    bad = torch.zeros(2, dtype=torch.int32, device="cuda")
    S = torch.cuda.current_stream().cuda_stream
    for rep in range(args.reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _capi.call("rti_ptm_fit_table_frames", d_f.data_ptr(), p, d_c.data_ptr(), n, 0, rows, cols,
                   axis.data_ptr(), g, coef.data_ptr(), p, table.data_ptr(), p, bad.data_ptr(), S)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        gb = (n + 48 + 4 * g * g) * p / 1e9
        print("rep %d: fused %.3f ms  %.0f GB/s  bad %s" % (rep, ms, gb / ms * 1e3, bad.tolist()))


if __name__ == "__main__":
    main()
