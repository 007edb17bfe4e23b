#!/usr/bin/env python
"""Per-kernel roofline measurements for every row of SURVEY.md section 8 that is not the
bench.py headline: gather, PTM fit (SoA / shared-light), fp64 grids, grids->table, relighting
(config 5: 4096x4096 coefficient map), linear RBF fit + evaluation (config 2 stand-in).

    python tools/bench_kernels.py [--out profiles/rNN_kernels.json] [--quick]

Each line: kernel, shape, ms (CUDA events on the launching stream, median of `reps` after 3
warm-ups, L2 flushed between repetitions by writing a 256 MB buffer when the working set is
small), algorithmic bytes or flops, achieved rate, fraction of the measured peak.
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


def hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class Timer:
    def __init__(self):
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def run(self, fn, reps=10, warmup=3, flush=True):
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize()
        times = []
        for _ in range(reps):
            if flush:
                self.flush.fill_(1)
            a = torch.cuda.Event(enable_timing=True)
            b = torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            times.append(a.elapsed_time(b))
        return float(np.median(times)), float(np.min(times))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    ap.add_argument("--quick", action="store_true")
    args = ap.parse_args()
    torch.cuda.set_device(0)
    _capi.load()
    peak, peak_src = hbm_peak()
    tm = Timer()
    S = torch.cuda.current_stream().cuda_stream
    out = []

    def emit(**kw):
        kw["peak_hbm_gbs"] = peak
        kw["peak_source"] = peak_src
        out.append(kw)
        print(json.dumps(kw), flush=True)

    def hbm_line(name, shape, ms, best, nbytes, **extra):
        gbs = nbytes / (ms * 1e-3) / 1e9
        emit(kernel=name, shape=shape, ms=ms, ms_best=best, algorithmic_bytes=nbytes, achieved_gbs=gbs,
             bound="hbm", frac=gbs / peak, **extra)

    # ---------------------------------------------------------------- gather (row a2)
    rows, cols, n = (256, 1024, 500) if not args.quick else (64, 256, 128)
    frames, cams = synthetic.random_stack(rows, cols, n, seed=0)
    d_f = torch.from_numpy(frames).cuda()
    d_c = torch.from_numpy(cams).cuda()
    p = rows * cols
    lx = torch.empty((p, n), dtype=torch.float32, device="cuda")
    ly = torch.empty_like(lx)
    it = torch.empty((p, n), dtype=torch.int32, device="cuda")
    ms, best = tm.run(lambda: _capi.call("rti_gather_soa", d_f.data_ptr(), p, d_c.data_ptr(), n, 0, rows, cols,
                                         lx.data_ptr(), ly.data_ptr(), it.data_ptr(), S), flush=False)
    hbm_line("gather_soa_kernel", [rows, cols, n], ms, best, 13 * p * n, units_per_s=p * n / (ms * 1e-3))

    # ---------------------------------------------------------------- PTM fit from SoA (row a4)
    coef = torch.empty((6, p), dtype=torch.float64, device="cuda")
    bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    ms, best = tm.run(lambda: _capi.call("rti_ptm_fit_soa", lx.data_ptr(), ly.data_ptr(), it.data_ptr(), p, n,
                                         coef.data_ptr(), p, bad.data_ptr(), S), flush=False)
    hbm_line("ptm_fit_soa_kernel", [p, n], ms, best, 12 * p * n + 48 * p, units_per_s=p * n / (ms * 1e-3))
    del lx, ly, it

    # ---------------------------------------------------------------- PTM fused fit (row a4), for reference
    ms, best = tm.run(lambda: _capi.call("rti_ptm_fit_frames", d_f.data_ptr(), p, d_c.data_ptr(), n, 0, rows, cols,
                                         coef.data_ptr(), p, bad.data_ptr(), S), flush=False)
    emit(kernel="ptm_fit_frames_kernel", shape=[rows, cols, n], ms=ms, ms_best=best, bound="fp64 pipe",
         units_per_s=p * n / (ms * 1e-3), fp64_slots_per_unit=50,
         est_fp64_tslots_per_s=50 * p * n / (ms * 1e-3) / 1e12,
         achieved_gbs=(p * n + 48 * p) / (ms * 1e-3) / 1e9)

    # ---------------------------------------------------------------- PTM shared-light fit
    pinv = torch.from_numpy(np.random.default_rng(0).normal(0, 1, (6, n))).cuda()
    ms, best = tm.run(lambda: _capi.call("rti_ptm_fit_shared", d_f.data_ptr(), p, pinv.data_ptr(), n, p,
                                         coef.data_ptr(), p, S), flush=False)
    # 6 DFMA per byte read: fp64-pipe bound (measured DFMA peak 18.5e12/s, profiles/r01_peak_fp64.jsonl)
    emit(kernel="ptm_fit_shared_kernel", shape=[p, n], ms=ms, ms_best=best, bound="fp64 pipe",
         units_per_s=p * n / (ms * 1e-3), dfma_per_s=6 * p * n / (ms * 1e-3),
         frac_of_dfma_peak=6 * p * n / (ms * 1e-3) / 18.5e12, achieved_gbs=(p * n + 48 * p) / (ms * 1e-3) / 1e9)
    del d_f

    # ---------------------------------------------------------------- ROI extraction (rows f2 / f4)
    fh, fw, nf, roi = (2160, 3840, 40, 2048) if not args.quick else (1080, 1920, 8, 400)
    bgr = torch.randint(0, 256, (nf, fh, fw, 3), dtype=torch.uint8, device="cuda")
    x0, y0 = (fw - roi) // 2 + 1, (fh - roi) // 2 + 1
    for hsv_flag, bpp in ((0, 1), (1, 3)):
        o = torch.empty((nf, roi, roi, bpp), dtype=torch.uint8, device="cuda")
        ms, best = tm.run(lambda: _capi.call("rti_roi_extract", bgr.data_ptr(), fh * fw * 3, nf, fh, fw, x0, y0, roi, roi,
                                             hsv_flag, o.data_ptr(), roi * roi * bpp, S))
        hbm_line("roi_extract_kernel<%s>" % ("hsv" if hsv_flag else "V"), [nf, roi, roi], ms, best,
                 (3 + bpp) * nf * roi * roi, frames_per_s=nf / (ms * 1e-3))
        del o
    del bgr

    # ---------------------------------------------------------------- fp64 grids + grids->table (rows a3, a6)
    axis = rti_b200.light_axis(0.02)
    d_axis = torch.from_numpy(axis).cuda()
    pg = 65536 if not args.quick else 4096
    grids = torch.empty((pg, 100, 100), dtype=torch.float64, device="cuda")
    ms, best = tm.run(lambda: _capi.call("rti_ptm_grids", coef.data_ptr(), p, pg, d_axis.data_ptr(), 100,
                                         grids.data_ptr(), S), flush=False)
    hbm_line("ptm_grids_kernel", [pg, 100, 100], ms, best, 8 * pg * 10000 + 48 * pg)
    table = torch.empty((100, 100, pg), dtype=torch.int32, device="cuda")
    ms, best = tm.run(lambda: _capi.call("rti_grids_to_table", grids.data_ptr(), pg, 100, table.data_ptr(), pg,
                                         bad.data_ptr(), S), flush=False)
    hbm_line("grids_to_table_kernel", [pg, 100, 100], ms, best, 12 * pg * 10000)
    del grids, table

    # ---------------------------------------------------------------- relighting, config 5 (row a8)
    R = 4096 if not args.quick else 1024
    P = R * R
    rng = np.random.default_rng(1)
    c64 = torch.from_numpy(rng.normal(0, 60, (6, P)) + np.array([0, 0, 0, 0, 0, 120.0])[:, None]).cuda()
    c32 = torch.empty((6, P), dtype=torch.float32, device="cuda")
    _capi.call("rti_coef_to_f32", c64.data_ptr(), 6 * P, c32.data_ptr(), S)
    hsv = torch.from_numpy(rng.integers(0, 180, (P, 3), dtype=np.uint8)).cuda()
    out_v = torch.empty(P, dtype=torch.uint8, device="cuda")
    out_bgr = torch.empty((P, 3), dtype=torch.uint8, device="cuda")
    dirs = [(float(a), float(b)) for a, b in rng.uniform(-1, 1, (1000, 2))]
    for name, fn_name, cmap, bpp in (("f64", "rti_relight_ptm", c64, 48), ("f32", "rti_relight_ptm_f32", c32, 24)):
        for hsv_t, o, io in ((None, out_v, 1), (hsv, out_bgr, 6)):
            state = {"i": 0}

            def one():
                lu, lv = dirs[state["i"] % 1000]
                state["i"] += 1
                _capi.call(fn_name, cmap.data_ptr(), P, P, R, lu, lv, hsv_t.data_ptr() if hsv_t is not None else None,
                           o.data_ptr(), S)
            # the coefficient map (403 / 201 MB) exceeds L2, no flush needed
            ms, best = tm.run(one, reps=50, flush=False)
            hbm_line("relight_kernel<%s,%s>" % (name, "hsv2bgr" if hsv_t is not None else "V"), [R, R], ms, best,
                     (bpp + io) * P, frames_per_s=1e3 / ms)
    del c64, c32, hsv, out_v, out_bgr

    # ---------------------------------------------------------------- linear RBF (row a5), config-2 stand-in
    for n_rbf, px in ((300, 148 * 16), (128, 148 * 32), (500, 148 * 4)) if not args.quick else ((128, 296),):
        rr = np.sqrt(rng.uniform(0, 0.95, (px, n_rbf)))
        th = rng.uniform(0, 2 * np.pi, (px, n_rbf))
        nlx = torch.from_numpy((rr * np.cos(th)).astype(np.float32)).cuda()
        nly = torch.from_numpy((rr * np.sin(th)).astype(np.float32)).cuda()
        nin = torch.from_numpy(rng.integers(0, 256, (px, n_rbf)).astype(np.int32)).cuda()
        w = torch.empty((px, n_rbf), dtype=torch.float64, device="cuda")
        lib = _capi.load()
        batch = 148 * 2
        nbytes = lib.rti_rbf_workspace_bytes(n_rbf, batch)
        work = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
        ms, best = tm.run(lambda: _capi.call("rti_rbf_fit_soa", nlx.data_ptr(), nly.data_ptr(), nin.data_ptr(), px,
                                             n_rbf, w.data_ptr(), work.data_ptr(), nbytes, batch, bad.data_ptr(), S),
                          reps=5, flush=False)
        flops = 2.0 * (n_rbf - 1) ** 3 / 3.0 * px
        emit(kernel="rbf_fit_kernel", shape=[px, n_rbf], ms=ms, ms_best=best, bound="fp64 pipe",
             pixels_per_s=px / (ms * 1e-3), cholesky_tflops=flops / (ms * 1e-3) / 1e12,
             mpixel_lights_per_s=px * n_rbf / (ms * 1e-3) / 1e6)
        tab = torch.empty((100, 100, px), dtype=torch.int32, device="cuda")
        ms, best = tm.run(lambda: _capi.call("rti_rbf_table", nlx.data_ptr(), nly.data_ptr(), w.data_ptr(), px, n_rbf,
                                             d_axis.data_ptr(), 100, tab.data_ptr(), px, bad.data_ptr(), S),
                          reps=5, flush=False)
        terms = px * 10000.0 * n_rbf
        emit(kernel="rbf_eval_kernel<8,table>", shape=[px, n_rbf, 100, 100], ms=ms, ms_best=best, bound="fp64 pipe",
             gterms_per_s=terms / (ms * 1e-3) / 1e9, pixels_per_s=px / (ms * 1e-3),
             mpixel_lights_per_s=px * n_rbf / (ms * 1e-3) / 1e6)
        del tab, work, w

    if args.out:
        with open(os.path.join(ROOT, args.out), "w") as f:
            json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
