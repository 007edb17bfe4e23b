#!/usr/bin/env python
"""Minimal driver for ncu captures / timings of the shared-light PTM contraction (rti_ptm_fit_shared).

    python tools/prof_shared_fit.py [--rows 256] [--cols 1024] [--n 500] [--reps 5]
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from rti_b200 import _capi  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=256)
    ap.add_argument("--cols", type=int, default=1024)
    ap.add_argument("--n", type=int, default=500)
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()
    torch.cuda.set_device(0)
    _capi.load()
    p, n = args.rows * args.cols, args.n
    frames = torch.randint(0, 256, (n, p), dtype=torch.uint8, device="cuda")
    pinv = torch.from_numpy(np.random.default_rng(0).normal(0, 1, (6, n))).cuda()
    coef = torch.empty((6, p), dtype=torch.float64, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    S = torch.cuda.current_stream().cuda_stream
    ts = []
    for _ in range(args.reps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        _capi.call("rti_ptm_fit_shared", frames.data_ptr(), p, pinv.data_ptr(), n, p, coef.data_ptr(), p, S)
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ms = min(ts)
    ref = (pinv @ frames[:, :4096].double())
    err = ((coef[:, :4096] - ref).abs().max() / ref.abs().max()).item()
    print("ptm_fit_shared %d px x %d: %.4f ms  %.0f GB/s  %.2f T DFMA/s (%.2f of 18.5)  rel err vs torch %.1e"
          % (p, n, ms, (p * n + 48 * p) / ms / 1e6, 6 * p * n / ms / 1e9, 6 * p * n / ms / 1e9 / 18.5, err))


if __name__ == "__main__":
    main()
