#!/usr/bin/env python
"""Minimal driver for ncu captures of the gather / PTM-fit kernels (one launch each after a warm-up).

    python tools/prof_ptm.py [--rows 256] [--cols 1024] [--n 500]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from rti_b200 import _capi, synthetic  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=256)
    ap.add_argument("--cols", type=int, default=1024)
    ap.add_argument("--n", type=int, default=500)
    args = ap.parse_args()
    torch.cuda.set_device(0)
    _capi.load()
    rows, cols, n = args.rows, args.cols, args.n
    frames, cams = synthetic.random_stack(rows, cols, n, seed=0)
    d_f, d_c = torch.from_numpy(frames).cuda(), torch.from_numpy(cams).cuda()
    p = rows * cols
    lx = torch.empty((p, n), dtype=torch.float32, device="cuda")
    ly = torch.empty_like(lx)
    it = torch.empty((p, n), dtype=torch.int32, device="cuda")
    coef = torch.empty((6, p), dtype=torch.float64, device="cuda")
    bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    S = torch.cuda.current_stream().cuda_stream
    for rep in range(2):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        ev[0].record()
        _capi.call("rti_gather_soa", d_f.data_ptr(), p, d_c.data_ptr(), n, 0, rows, cols,
                   lx.data_ptr(), ly.data_ptr(), it.data_ptr(), S)
        ev[1].record()
        _capi.call("rti_ptm_fit_soa", lx.data_ptr(), ly.data_ptr(), it.data_ptr(), p, n,
                   coef.data_ptr(), p, bad.data_ptr(), S)
        ev[2].record()
        _capi.call("rti_ptm_fit_frames", d_f.data_ptr(), p, d_c.data_ptr(), n, 0, rows, cols,
                   coef.data_ptr(), p, bad.data_ptr(), S)
        ev[3].record()
        torch.cuda.synchronize()
        print("rep %d: gather %.3f ms  fit_soa %.3f ms  fit_frames %.3f ms"
              % (rep, ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2]), ev[2].elapsed_time(ev[3])))


if __name__ == "__main__":
    main()
