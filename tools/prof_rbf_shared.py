#!/usr/bin/env python
"""Timing / ncu driver of the shared-light linear-RBF contraction (tcgen05):

    python tools/prof_rbf_shared.py [--n 500] [--pixels 1048576] [--grid 100] [--reps 5] [--json out.json]

Prints the operator build time (per-pixel fp64 kernels on n pseudo-pixels + fp16 hi/lo split),
the uint8 -> fp16 pack time, and the table contraction: ms, executed tensor TFLOP/s
(2 MMAs per k-step on the padded shape), useful TFLOP/s (2.G^2.n.P), table GB/s.
"""
import argparse
import json
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
    ap.add_argument("--n", type=int, default=500)
    ap.add_argument("--pixels", type=int, default=1024 * 1024)
    ap.add_argument("--grid", type=int, default=100)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    torch.cuda.set_device(0)
    lib = _capi.load()
    n, p, g = args.n, args.pixels, args.grid
    rows = g * g
    rng = np.random.default_rng(0)
    rr = np.sqrt(rng.uniform(0, 0.95, n))
    th = rng.uniform(0, 2 * np.pi, n)
    lx = torch.from_numpy((rr * np.cos(th)).astype(np.float32)).cuda()
    ly = torch.from_numpy((rr * np.sin(th)).astype(np.float32)).cuda()
    axis = torch.from_numpy(np.linspace(-1.0, 1.0, g, endpoint=False)).cuda()
    frames = torch.randint(0, 256, (n, p), dtype=torch.uint8, device="cuda")
    elems = lib.rti_rbf_shared_operator_elems(rows, n)
    hi = torch.empty(elems, dtype=torch.int16, device="cuda")
    lo = torch.empty(elems, dtype=torch.int16, device="cuda")
    inv = torch.empty(rows, dtype=torch.float32, device="cuda")
    bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    packed = torch.empty(lib.rti_rbf_shared_packed_elems(n, p), dtype=torch.int16, device="cuda")
    table = torch.empty((rows, p), dtype=torch.int32, device="cuda")
    S = torch.cuda.current_stream().cuda_stream

    def timed(fn, reps):
        out = []
        for _ in range(reps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            out.append(a.elapsed_time(b))
        return out

    t_op = timed(lambda: _capi.call("rti_rbf_shared_operator", lx.data_ptr(), ly.data_ptr(), n, axis.data_ptr(), g, 0,
                                    hi.data_ptr(), lo.data_ptr(), inv.data_ptr(), bad.data_ptr(), S), 2)
    t_pack = timed(lambda: _capi.call("rti_rbf_shared_pack_frames", frames.data_ptr(), p, n, p, packed.data_ptr(), S),
                   max(2, args.reps))
    sys.path.insert(0, ROOT)
    from bench import ClockSampler
    sampler = ClockSampler(0)
    sampler.start()
    t_gemm = timed(lambda: _capi.call("rti_rbf_shared_table", packed.data_ptr(), n, p, hi.data_ptr(), lo.data_ptr(),
                                      inv.data_ptr(), rows, table.data_ptr(), p, bad.data_ptr(), S), args.reps + 1)[1:]
    clocks = sampler.stop()
    ms = float(np.median(t_gemm))
    n_pad = -(-n // 64) * 64
    r_pad = -(-rows // 128) * 128
    p_pad = -(-p // 256) * 256
    executed = 2.0 * 2.0 * r_pad * n_pad * p_pad
    useful = 2.0 * rows * n * p
    res = {
        "kernel": "rbf_shared_gemm_kernel<tma_store>", "shape": {"lights": n, "pixels": p, "grid": [g, g]},
        "operator_build_ms": float(min(t_op)), "pack_frames_ms": float(np.median(t_pack)),
        "ms": ms, "ms_best": float(min(t_gemm)),
        "executed_tflops_fp16": executed / ms / 1e9, "useful_tflops": useful / ms / 1e9,
        "table_gbs": rows * p * 4 / ms / 1e6, "mpixel_lights_per_s": p * n / ms / 1e3,
        "bad": int(bad.item()), "clocks": clocks, "all_ms": [round(t, 3) for t in t_gemm],
    }
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        res["peak_tflops_bf16_sustained"] = peaks.get("bf16_tflops_sustained")
        res["frac_of_tensor_peak"] = res["executed_tflops_fp16"] / peaks["bf16_tflops_sustained"]
    except Exception:
        pass
    print(json.dumps(res))
    if args.json:
        with open(args.json, "w") as f:
            json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
