#!/usr/bin/env python
"""BASELINE.json configs[3] on ONE GPU: synthetic 4096x4096 ROI x 1000 light samples, PTM and
linear RBF, 100x100 light grid.

The int32 table of this configuration is 671 GB, so it cannot be resident: the ROI is processed in
row slabs of 256 rows (1 M pixels); every slab's 42 GB table slab is written into ONE reused device
buffer (far larger than L2), exactly like a streaming writer would consume it.  The frames (16.8 GB)
and the fp64 coefficient map (805 MB) stay resident.  Reported per form:

  ptm                 fused gather + fit + table kernel over all 16 slabs (complete run)
  rbf_shared_lights   pack + tcgen05 contraction over all 16 slabs (complete run)
  rbf_per_pixel       fp64 fit + evaluation on a pixel SAMPLE, extrapolated linearly in pixels
                      (every pixel is independent and costs the same)

    python tools/bench_config4.py [--json profiles/rNN_config4.json] [--slab-rows 256]
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
from rti_b200 import _capi, synthetic  # noqa: E402

R, N, G = 4096, 1000, 100


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--json", default=None)
    ap.add_argument("--slab-rows", type=int, default=256)
    ap.add_argument("--rbf-sample", type=int, default=148 * 8)
    args = ap.parse_args()
    torch.cuda.set_device(0)
    lib = _capi.load()
    S = torch.cuda.current_stream().cuda_stream
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    P = R * R
    slab_rows = args.slab_rows
    slab_px = slab_rows * R
    n_slabs = R // slab_rows
    gen = torch.Generator(device="cuda").manual_seed(0)
    frames = torch.empty((N, R, R), dtype=torch.uint8, device="cuda")
    for i in range(0, N, 50):                                   # uniform-random uint8 ROI stack, 16.8 GB
        frames[i:i + 50] = torch.randint(0, 256, (min(50, N - i), R, R), dtype=torch.uint8, device="cuda",
                                         generator=gen)
    cams = synthetic.camera_positions(R, R, N, np.random.default_rng(0))
    d_cams = torch.from_numpy(cams).cuda()
    axis = rti_b200.light_axis(0.02)
    d_axis = torch.from_numpy(axis).cuda()
    coef = torch.empty((6, P), dtype=torch.float64, device="cuda")
    table = torch.empty((G * G, slab_px), dtype=torch.int32, device="cuda")       # reused slab buffer, 42 GB
    bad = torch.zeros(2, dtype=torch.int32, device="cuda")
    out = {"config": "synthetic 4096x4096 ROI x 1000 lights, 100x100 grid, 1 GPU, %d slabs of %d rows" % (n_slabs, slab_rows),
           "units_pixel_lights": P * N}

    def run_ptm():
        for s in range(n_slabs):
            _capi.call("rti_ptm_fit_table_frames", frames.data_ptr() + s * slab_px, P, d_cams.data_ptr(), N,
                       s * slab_rows, slab_rows, R, d_axis.data_ptr(), G, coef.data_ptr() + 8 * s * slab_px, P,
                       table.data_ptr(), slab_px, bad.data_ptr(), S)

    def timed(fn, reps=2):
        fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(reps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        return min(ts)

    ms = timed(run_ptm)
    alg = (N + 48 + 4 * G * G) * P
    out["ptm"] = {"ms": ms, "mpixel_lights_per_s": P * N / ms / 1e3, "algorithmic_bytes": alg,
                  "achieved_gbs": alg / ms / 1e6, "hbm_peak_gbs": hbm_peak, "frac": alg / ms / 1e6 / hbm_peak,
                  "bad": bad.cpu().tolist(), "what": "fused gather + fit + int32 table, complete run"}
    print(json.dumps({"ptm": out["ptm"]}), flush=True)

    # ---- shared-light RBF: nodes = the light vectors seen from the ROI centre
    c = np.array([R / 2.0, R / 2.0, 0.0])
    l = (cams - c) / np.linalg.norm(cams - c, axis=1, keepdims=True)
    d_lx = torch.from_numpy(l[:, 0].astype(np.float32)).cuda()
    d_ly = torch.from_numpy(l[:, 1].astype(np.float32)).cuda()
    rows = G * G
    elems = lib.rti_rbf_shared_operator_elems(rows, N)
    hi = torch.empty(elems, dtype=torch.int16, device="cuda")
    lo = torch.empty(elems, dtype=torch.int16, device="cuda")
    inv = torch.empty(rows, dtype=torch.float32, device="cuda")
    packed = torch.empty(lib.rti_rbf_shared_packed_elems(N, slab_px), dtype=torch.int16, device="cuda")
    sbad = torch.zeros(1, dtype=torch.int32, device="cuda")
    op_ms = timed(lambda: _capi.call("rti_rbf_shared_operator", d_lx.data_ptr(), d_ly.data_ptr(), N, d_axis.data_ptr(),
                                     G, 0, hi.data_ptr(), lo.data_ptr(), inv.data_ptr(), sbad.data_ptr(), S), reps=1)

    def run_shared():
        for s in range(n_slabs):
            _capi.call("rti_rbf_shared_pack_frames", frames.data_ptr() + s * slab_px, P, N, slab_px,
                       packed.data_ptr(), S)
            _capi.call("rti_rbf_shared_table", packed.data_ptr(), N, slab_px, hi.data_ptr(), lo.data_ptr(),
                       inv.data_ptr(), rows, table.data_ptr(), slab_px, sbad.data_ptr(), S)

    ms = timed(run_shared, reps=1)
    n_pad, r_pad = -(-N // 64) * 64, -(-rows // 128) * 128
    executed = 4.0 * r_pad * n_pad * P / ms / 1e9
    out["rbf_shared_lights"] = {
        "ms": ms, "operator_build_ms": op_ms, "mpixel_lights_per_s": P * N / ms / 1e3,
        "executed_tflops_fp16": executed, "useful_tflops": 2.0 * rows * N * P / ms / 1e9,
        "peak_tflops_burst": peaks.get("bf16_tflops"), "peak_tflops_sustained": peaks.get("bf16_tflops_sustained"),
        "frac_of_sustained": executed / peaks["bf16_tflops_sustained"] if peaks else None,
        "bad": int(sbad.item()), "what": "pack + tcgen05 contraction into the int32 table, complete run"}
    print(json.dumps({"rbf_shared_lights": out["rbf_shared_lights"]}), flush=True)
    del packed, hi, lo

    # ---- per-pixel RBF on a sample of one ROI row segment
    px = args.rbf_sample
    lx = torch.empty((px, N), dtype=torch.float32, device="cuda")
    ly = torch.empty_like(lx)
    inten = torch.empty((px, N), dtype=torch.int32, device="cuda")
    _capi.call("rti_gather_soa", frames.data_ptr() + 2048 * R, P, d_cams.data_ptr(), N, 2048, 1, px,
               lx.data_ptr(), ly.data_ptr(), inten.data_ptr(), S)
    w = torch.empty((px, N), dtype=torch.float64, device="cuda")
    batch = 148 * 2
    nbytes = lib.rti_rbf_workspace_bytes(N, batch)
    work = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    tab = torch.empty((G, G, px), dtype=torch.int32, device="cuda")
    fit_ms = timed(lambda: _capi.call("rti_rbf_fit_soa", lx.data_ptr(), ly.data_ptr(), inten.data_ptr(), px, N,
                                      w.data_ptr(), work.data_ptr(), nbytes, batch, sbad.data_ptr(), S))
    ev_ms = timed(lambda: _capi.call("rti_rbf_table", lx.data_ptr(), ly.data_ptr(), w.data_ptr(), px, N,
                                     d_axis.data_ptr(), G, tab.data_ptr(), px, sbad.data_ptr(), S))
    scale = P / px
    out["rbf_per_pixel"] = {
        "sample_pixels": px, "fit_ms": fit_ms, "table_ms": ev_ms,
        "mpixel_lights_per_s": px * N / (fit_ms + ev_ms) / 1e3,
        "fit_fp64_tflops": (N ** 3 / 3.0) * 2 * px / fit_ms / 1e9,
        "extrapolated_full_roi_s": (fit_ms + ev_ms) * scale / 1e3,
        "what": "fp64 DMMA Cholesky + fp64 grid evaluation on a pixel sample; sampled, extrapolated",
        "bad": int(sbad.item())}
    print(json.dumps({"rbf_per_pixel": out["rbf_per_pixel"]}), flush=True)
    if args.json:
        with open(args.json, "w") as f:
            json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
