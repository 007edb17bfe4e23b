#!/usr/bin/env python
"""Times the tile-preconditioned linear-RBF fit (rti_rbf_fit_pcg) next to the direct fp64 fit on a
rows x cols window of a 1024^2 ROI (the bench geometry): ms, us/pixel, iteration statistics.

    python tools/prof_rbf_pcg.py [--n 500] [--rows 64] [--cols 256] [--max-iters 60] [--direct]
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import rti_b200  # noqa: E402
from rti_b200 import _capi, synthetic  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=500)
    ap.add_argument("--rows", type=int, default=64)
    ap.add_argument("--cols", type=int, default=256)
    ap.add_argument("--roi", type=int, default=1024)
    ap.add_argument("--max-iters", type=int, default=60)
    ap.add_argument("--tol", type=float, default=0.01)
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--direct", action="store_true")
    ap.add_argument("--lambert", action="store_true")
    args = ap.parse_args()
    torch.cuda.set_device(0)
    lib = _capi.load()
    S = torch.cuda.current_stream().cuda_stream
    rows, cols, n = args.rows, args.cols, args.n
    px = rows * cols
    if args.lambert:
        frames, _ = synthetic.lambertian_stack(rows, cols, n, seed=1, row0=400, rows_total=args.roi)
    else:
        frames = np.random.default_rng(1).integers(0, 256, (n, rows, cols), dtype=np.uint8)
    cams = synthetic.camera_positions(args.roi, args.roi, n, np.random.default_rng(0))
    d_frames = torch.from_numpy(frames).cuda()
    d_cams = torch.from_numpy(cams).cuda()
    lx = torch.empty((px, n), dtype=torch.float32, device="cuda")
    ly = torch.empty_like(lx)
    inten = torch.empty((px, n), dtype=torch.int32, device="cuda")
    _capi.call("rti_gather_soa", d_frames.data_ptr(), px, d_cams.data_ptr(), n, 400, rows, cols,
               lx.data_ptr(), ly.data_ptr(), inten.data_ptr(), S)
    w = torch.empty((px, n), dtype=torch.float64, device="cuda")
    bad = torch.zeros(2, dtype=torch.int32, device="cuda")
    iters = torch.zeros(px, dtype=torch.int32, device="cuda")
    tiles = ((rows + 15) // 16) * ((cols + 15) // 16)
    nbytes = lib.rti_rbf_pcg_workspace_bytes(n, tiles)
    work = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    for _ in range(args.reps):
        bad.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        _capi.call("rti_rbf_fit_pcg", lx.data_ptr(), ly.data_ptr(), inten.data_ptr(), rows, cols, n, w.data_ptr(),
                   work.data_ptr(), nbytes, args.max_iters, args.tol, iters.data_ptr(), bad.data_ptr(), S)
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b)
        it = iters.cpu().numpy()
        print("pcg   n=%d %dx%d: %.2f ms = %.3f us/pixel (%.0f Mpixel.lights/s); steps mean %.1f p95 %d max %d; "
              "singular tiles %d, unconverged %d; workspace %.2f GB"
              % (n, rows, cols, ms, ms * 1e3 / px, px * n / ms / 1e3, it.mean(), np.percentile(it, 95), it.max(),
                 int(bad[0].item()), int(bad[1].item()), nbytes / 1e9))
    if args.direct:
        w2 = torch.empty((px, n), dtype=torch.float64, device="cuda")
        batch = 296
        nb = lib.rti_rbf_workspace_bytes(n, batch)
        wk = torch.empty(nb, dtype=torch.uint8, device="cuda")
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        _capi.call("rti_rbf_fit_soa", lx.data_ptr(), ly.data_ptr(), inten.data_ptr(), px, n, w2.data_ptr(),
                   wk.data_ptr(), nb, batch, bad.data_ptr(), S)
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b)
        rel = ((w - w2).abs().amax(dim=1) / w2.abs().amax(dim=1)).max().item()
        print("direct n=%d: %.2f ms = %.3f us/pixel; max rel weight difference pcg vs direct %.2e" % (n, ms, ms * 1e3 / px, rel))


if __name__ == "__main__":
    main()
