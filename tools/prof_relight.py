#!/usr/bin/env python
"""Minimal driver for ncu captures of the relight kernels (config 5: one launch of each form on a
4096 x 4096 coefficient map: f64 V, f64 HSV->BGR, f32 V, f32 HSV->BGR, then the table-lookup form).

    python tools/prof_relight.py [--size 4096]
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
    ap.add_argument("--size", type=int, default=4096)
    args = ap.parse_args()
    torch.cuda.set_device(0)
    _capi.load()
    R = args.size
    P = R * R
    S = torch.cuda.current_stream().cuda_stream
    rng = np.random.default_rng(1)
    c64 = torch.from_numpy(rng.normal(0, 60, (6, P)) + np.array([0, 0, 0, 0, 0, 120.0])[:, None]).cuda()
    c32 = torch.empty((6, P), dtype=torch.float32, device="cuda")
    _capi.call("rti_coef_to_f32", c64.data_ptr(), 6 * P, c32.data_ptr(), S)
    hsv = torch.from_numpy(rng.integers(0, 180, (P, 3), dtype=np.uint8)).cuda()
    out_v = torch.empty(P, dtype=torch.uint8, device="cuda")
    out_bgr = torch.empty((P, 3), dtype=torch.uint8, device="cuda")
    image = torch.from_numpy(rng.integers(-40, 300, P).astype(np.int32)).cuda()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
    ev[0].record()
    _capi.call("rti_relight_ptm", c64.data_ptr(), P, P, R, 0.31, -0.42, None, out_v.data_ptr(), S)
    ev[1].record()
    _capi.call("rti_relight_ptm", c64.data_ptr(), P, P, R, 0.31, -0.42, hsv.data_ptr(), out_bgr.data_ptr(), S)
    ev[2].record()
    _capi.call("rti_relight_ptm_f32", c32.data_ptr(), P, P, R, 0.31, -0.42, None, out_v.data_ptr(), S)
    ev[3].record()
    _capi.call("rti_relight_ptm_f32", c32.data_ptr(), P, P, R, 0.31, -0.42, hsv.data_ptr(), out_bgr.data_ptr(), S)
    ev[4].record()
    _capi.call("rti_relight_from_image", image.data_ptr(), P, R, hsv.data_ptr(), out_bgr.data_ptr(), S)
    ev[5].record()
    torch.cuda.synchronize()
    print("f64 V %.3f ms  f64 BGR %.3f ms  f32 V %.3f ms  f32 BGR %.3f ms  table BGR %.3f ms (first launches)"
          % tuple(ev[i].elapsed_time(ev[i + 1]) for i in range(5)))


if __name__ == "__main__":
    main()
