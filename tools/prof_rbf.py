#!/usr/bin/env python
"""Minimal driver for ncu captures of the linear-RBF kernels: one fit + one table launch.

    python tools/prof_rbf.py [--n 300] [--pixels 2368] [--reps 1]
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import rti_b200  # noqa: E402
from rti_b200 import _capi  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=300)
    ap.add_argument("--pixels", type=int, default=148 * 16)
    ap.add_argument("--reps", type=int, default=1)
    ap.add_argument("--batch", type=int, default=148 * 2)
    ap.add_argument("--no-eval", action="store_true")
    ap.add_argument("--fast", action="store_true", help="also time rti_rbf_table_fast (float32 distances)")
    args = ap.parse_args()
    torch.cuda.set_device(0)
    lib = _capi.load()
    rng = np.random.default_rng(1)
    px, n = args.pixels, args.n
    rr = np.sqrt(rng.uniform(0, 0.95, (px, n)))
    th = rng.uniform(0, 2 * np.pi, (px, n))
    lx = torch.from_numpy((rr * np.cos(th)).astype(np.float32)).cuda()
    ly = torch.from_numpy((rr * np.sin(th)).astype(np.float32)).cuda()
    inten = torch.from_numpy(rng.integers(0, 256, (px, n)).astype(np.int32)).cuda()
    w = torch.empty((px, n), dtype=torch.float64, device="cuda")
    bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    nbytes = lib.rti_rbf_workspace_bytes(n, args.batch)
    work = torch.empty(max(nbytes, 8), dtype=torch.uint8, device="cuda")
    axis = torch.from_numpy(rti_b200.light_axis(0.02)).cuda()
    tab = torch.empty((100, 100, px), dtype=torch.int32, device="cuda")
    S = torch.cuda.current_stream().cuda_stream
    for _ in range(args.reps):
        a, b, c = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        a.record()
        _capi.call("rti_rbf_fit_soa", lx.data_ptr(), ly.data_ptr(), inten.data_ptr(), px, n, w.data_ptr(),
                   work.data_ptr(), nbytes, args.batch, bad.data_ptr(), S)
        b.record()
        if not args.no_eval:
            _capi.call("rti_rbf_table", lx.data_ptr(), ly.data_ptr(), w.data_ptr(), px, n, axis.data_ptr(), 100,
                       tab.data_ptr(), px, bad.data_ptr(), S)
        c.record()
        torch.cuda.synchronize()
        print("n=%d pixels=%d fit %.3f ms (%.0f px/s)  table %.3f ms  bad=%d"
              % (n, px, a.elapsed_time(b), px / a.elapsed_time(b) * 1e3, b.elapsed_time(c), int(bad.item())))
        if args.fast:
            d, e = (torch.cuda.Event(enable_timing=True) for _ in range(2))
            d.record()
            _capi.call("rti_rbf_table_fast", lx.data_ptr(), ly.data_ptr(), w.data_ptr(), px, n, axis.data_ptr(), 100,
                       tab.data_ptr(), px, bad.data_ptr(), S)
            e.record()
            torch.cuda.synchronize()
            ms = d.elapsed_time(e)
            sm_clk = 148 * 16 * 1.965e9
            print("   table_fast %.3f ms: %.3f us/pixel, %.2f Tsqrt/s = %.2f of the SFU peak (148 SM x 16/clk x 1.965 GHz)"
                  % (ms, ms * 1e3 / px, 1e4 * n * px / ms / 1e9, 1e4 * n * px / (ms * 1e-3) / sm_clk))


if __name__ == "__main__":
    main()
