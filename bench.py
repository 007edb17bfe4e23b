#!/usr/bin/env python
"""Benchmark of the per-pixel PTM fit + light-grid evaluation path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (BASELINE.json configs[2], the largest single-GPU configuration): synthetic
1024x1024 ROI x 500 light samples, PTM fit (fused from uint8 frames, per-pixel light
vectors) + evaluation on the 100x100 light grid into the reference's int32 table
layout.  With N > 1 (torchrun) every rank owns a 1024-row slab of a (1024*N) x 1024
ROI: weak scaling, no collective in the timed region.

One step = one launch of the fused gather+fit+table kernel over the rank's slab.  `value` is timed on the
device (CUDA events, inputs resident in HBM); `e2e` is the same work through the
host-buffer C-ABI call rti_ptm_compute_host (pinned host frames in, int32 table out).
Prints ONE JSON line (rank 0).
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ROWS, COLS, N_LIGHTS, G = 1024, 1024, 500, 100
METRIC = "Mpixel.lights/s fit+grid-eval (PTM, 100x100 light grid, int32 table)"
UNIT = "Mpixel.lights/s"
WORKLOAD = "synthetic 1024x1024 ROI x 500 light samples, PTM fit + 100x100 light grid, per GPU"


def read_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md: 6.65 TB/s)"


def read_traffic():
    """dram bytes per launch of the table kernel from the committed ncu --set full capture."""
    try:
        with open(os.path.join(ROOT, "profiles", "fused_kernel_traffic.json")) as f:
            d = json.load(f)
        if d.get("rows") == ROWS and d.get("cols") == COLS and d.get("g") == G:
            return float(d["dram_bytes_per_launch"])
    except Exception:
        pass
    return None


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.samples, self.reasons, self.max_mhz, self._stop_evt = [], set(), None, threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(nv, "nvmlClocksThrottleReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake",
        }
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop_evt.wait(0.005)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# ---------------------------------------------------------------------- CPU reference arm
REF_DIR = os.path.join(ROOT, "baseline", "_ref")


def make_config(world):
    """The `config` object of BOTH arms' JSON lines (key-identical so the driver can compare them)."""
    return {"workload": WORKLOAD, "rows_per_gpu": ROWS, "cols": COLS, "lights": N_LIGHTS, "grid": [G, G],
            "parallelism": "row-tile x%d, no collective" % world,
            "l2": "inputs (524 MB) and the 42 GB table written per step exceed the 126 MB L2"}


def _reference_module():
    """The UNMODIFIED reference analysis.py when a copy sits in baseline/_ref (git-ignored, never
    committed); matplotlib is stubbed because analysis.py:14 imports it for debug plots only."""
    if not os.path.isfile(os.path.join(REF_DIR, "analysis.py")):
        return None
    import types
    mpl, plt = types.ModuleType("matplotlib"), types.ModuleType("matplotlib.pyplot")
    mpl.pyplot = plt
    sys.modules.setdefault("matplotlib", mpl)
    sys.modules.setdefault("matplotlib.pyplot", plt)
    sys.path.insert(0, REF_DIR)
    cwd = os.getcwd()
    os.chdir(REF_DIR)
    try:
        import analysis
    except Exception:
        return None
    finally:
        os.chdir(cwd)
    return analysis


def _cpu_worker(args):
    """The per-pixel path for `count` pixels on one core.  mode "ptm": full-SVD PTM fit
    (analysis.py:295-298), grid evaluation (:301-315), int32 truncation (:404-407); mode "rbf":
    scipy Rbf('linear') fit + grid evaluation (analysis.py:249-260) + truncation.  Runs the
    unmodified reference functions when baseline/_ref holds them, else the oracle port."""
    seed, count, n, mode = args
    os.environ["OMP_NUM_THREADS"] = "1"
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import rti_oracle as orc
    ref = _reference_module()
    rng = np.random.default_rng(seed)
    lx = rng.uniform(-0.9, 0.9, (count, n)).astype(np.float32)
    ly = rng.uniform(-0.9, 0.9, (count, n)).astype(np.float32)
    inten = rng.integers(0, 256, (count, n)).astype(np.int32)
    ax = orc.light_axis()
    xi, yi = np.meshgrid(ax, ax, indexing="ij")
    t0 = time.perf_counter()
    acc = 0
    for p in range(count):
        if ref is not None:
            if mode == "ptm":
                grid = ref._interpolate_PTM(lx[p], ly[p], ax, inten[p])
            else:
                grid = ref._interpolate_RBF(lx[p], ly[p], xi, yi, inten[p])
            grid = np.asarray(grid)
        elif mode == "ptm":
            grid = orc.ptm_grid(orc.ptm_fit(lx[p], ly[p], inten[p]), ax)
        else:
            grid = orc.rbf_grid(lx[p], ly[p], orc.rbf_fit(lx[p], ly[p], inten[p]), ax)
        acc += int(orc.table_from_grids(grid[None, None])[0, 0, 0, 0])
    return time.perf_counter() - t0, acc, ref is not None


def cpu_reference(pixels_per_core, steps=1, warmup=0, mode="ptm", cores=None):
    """Times the CPU path on `cores` host cores (default: all); returns a cpu_baseline dict + ms/step."""
    import multiprocessing as mp
    try:
        import threadpoolctl
        threadpoolctl.threadpool_limits(1)
    except Exception:
        pass
    cores = len(os.sched_getaffinity(0)) if cores is None else cores
    ctx = mp.get_context("fork")
    with ctx.Pool(cores) as pool:
        kind = pool.map(_cpu_worker, [(i, 1, N_LIGHTS, mode) for i in range(cores)])[0][2]    # warm the workers
        for _ in range(warmup):
            pool.map(_cpu_worker, [(i, pixels_per_core, N_LIGHTS, mode) for i in range(cores)])
        t0 = time.perf_counter()
        for s in range(max(steps, 1)):
            pool.map(_cpu_worker, [(1000 * s + i, pixels_per_core, N_LIGHTS, mode) for i in range(cores)])
        dt = (time.perf_counter() - t0) / max(steps, 1)
    pixels = pixels_per_core * cores
    value = pixels * N_LIGHTS / dt / 1e6
    what = {"ptm": "full-SVD PTM fit + 100x100 grid eval + int32 truncation",
            "rbf": "linear-RBF solve(A, d) + cdist grid eval + int32 truncation"}[mode]
    who = "the unmodified reference functions (baseline/_ref)" if kind else "oracle port of"
    sample = ("%d random pixels x %d lights per step (%d per core), %s %s; extrapolates linearly in pixels"
              % (pixels, N_LIGHTS, pixels_per_core, who, what))
    out = {"value": value, "unit": UNIT, "cores": cores, "kind": "reference" if kind else "port", "sample": sample}
    if not kind:
        # the port against the UNMODIFIED reference functions, measured where the reference is mounted (the build
        # container, one core; tools/time_reference_cpu.py): how much slower the reference as written is
        try:
            with open(os.path.join(ROOT, "profiles", "r02_reference_vs_port_cpu.json")) as f:
                r = json.load(f)[mode]
            out["port_over_unmodified_reference"] = {
                "ratio": r["port_over_reference"], "where": "build container, 1 core, N=%d" % N_LIGHTS,
                "reference_ms_per_pixel": r["reference_ms_per_pixel"], "port_ms_per_pixel": r["port_ms_per_pixel"]}
        except Exception:
            pass
    return out, dt * 1e3


def run_reference(args, rank):
    if rank != 0:
        return
    cpu, ms = cpu_reference(args.cpu_pixels_per_core, steps=args.steps, warmup=min(args.warmup, 1))
    one, _ = cpu_reference(max(args.cpu_pixels_per_core // 4, 1), cores=1)
    cpu["one_core_value"] = one["value"]
    rbf, _ = cpu_reference(max(args.cpu_pixels_per_core // 12, 1), mode="rbf")
    line = {
        "impl": "reference", "metric": METRIC, "value": cpu["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": make_config(args.gpus),
        "cpu_baseline": cpu,
        "e2e": {"value": cpu["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "secondary": {"rbf_per_pixel_lights": {"cpu_baseline": rbf}},
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------- linear-RBF side measurements
def secondary_rbf(torch, lib, _capi, d_frames, d_cams, d_axis, p):
    """The two linear-RBF forms of the same workload on bounded pixel samples (not part of `value`):
    per-pixel lights (the reference's semantics: fp64 DMMA Cholesky + fp64 grid evaluation) on 2 ROI
    rows, and shared lights (one tcgen05 contraction) on 256 ROI rows.  Device-timed, best of 3."""
    S = torch.cuda.current_stream().cuda_stream

    def best(fn, reps=3):
        fn()
        torch.cuda.synchronize()
        out = []
        for _ in range(reps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            out.append(a.elapsed_time(b))
        return min(out)

    res = {}
    bad = torch.zeros(2, dtype=torch.int32, device="cuda")
    # square roots per second: 16 per clock per SM = 4.65e12 nominal; this repo's microbenchmark
    # (tools/peak_sfu.cu, profiles/r02_peak_sfu.jsonl) sustains 4.62e12 with MUFU alone and 4.14e12 with the
    # evaluation kernel's instruction mix on register operands
    sm_peak = 4.62e12
    # ---- per-pixel lights (the reference's semantics), default "fast" path: rows [0, 256) of the slab
    rows_s = 256
    px = rows_s * COLS
    lx = torch.empty((px, N_LIGHTS), dtype=torch.float32, device="cuda")
    ly = torch.empty_like(lx)
    inten = torch.empty((px, N_LIGHTS), dtype=torch.int32, device="cuda")
    _capi.call("rti_gather_soa", d_frames.data_ptr(), p, d_cams.data_ptr(), N_LIGHTS, 0, rows_s, COLS,
               lx.data_ptr(), ly.data_ptr(), inten.data_ptr(), S)
    from rti_b200 import analysis
    tab = torch.empty((G * G, px), dtype=torch.int32, device="cuda")
    max_iters, tol = 32, 0.01
    state = {}

    def pcg():
        # rti_rbf_fit_pcg + direct fp64 refit of the pixels it reports (not converged / weights too large for
        # float32 distances), as the drop-in runs it
        hbad = torch.zeros(1, dtype=torch.int32, device="cuda")
        state["w"], state["info"] = analysis.rbf_fit_device_pcg(lx, ly, inten, rows_s, COLS, hbad, max_iters=max_iters,
                                                              tol=tol)
        state["singular"] = hbad

    fit_ms = best(pcg, reps=2)
    w, info = state["w"], state["info"]
    singular_tiles, unconverged = info["singular_tiles"], info["refit"]
    bad.zero_()
    ev_ms = best(lambda: analysis.rbf_table_device(lx, ly, w, d_axis, G, tab, bad, fp64_idx=info["fp64_idx"]), reps=2)
    tab = tab.view(G, G, px)
    # sampled oracle check (scipy's algorithm in fp64: solve(A, d), cdist(grid, nodes) @ w)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import rti_oracle as orc
    ax = orc.light_axis()
    rng = np.random.default_rng(5)
    worst, flips = 0.0, 0
    sample_px = [int(v) for v in rng.integers(0, px, 6)]
    if info["fp64_idx"] is not None and info["fp64_idx"].numel():
        sample_px += [int(v) for v in info["fp64_idx"][:2].tolist()]        # and two of the fp64-routed pixels
    for q in sample_px:
        q = int(q)
        l1, l2, dd = lx[q].cpu().numpy(), ly[q].cpu().numpy(), inten[q].cpu().numpy()
        ref = orc.rbf_grid(l1, l2, orc.rbf_fit(l1, l2, dd), ax)
        got = orc.rbf_grid(l1, l2, w[q].cpu().numpy(), ax)
        worst = max(worst, float(np.max(np.abs(got - ref)) / np.max(np.abs(ref))))
        t = tab[:, :, q].cpu().numpy()
        d = t != np.trunc(ref).astype(np.int32)
        flips += int(np.count_nonzero(d & (np.abs(ref - np.rint(ref)) > 1e-3 * np.max(np.abs(ref)))))
    res["rbf_per_pixel_lights"] = {
        "sample": "%d pixels (%d ROI rows) x %d lights, fit + 100x100 table" % (px, rows_s, N_LIGHTS),
        "path": "tile-preconditioned CG fit (rti_rbf_fit_pcg: fp64 DMMA tile inverse, float32 distances + fp64 sums, "
                "batched tcgen05 preconditioner) + float32-distance table (rti_rbf_table_fast); the reported pixels "
                "through rti_rbf_fit_soa + rti_rbf_table (fp64), inside the timed calls",
        "fit_ms": fit_ms, "table_ms": ev_ms, "value": px * N_LIGHTS / ((fit_ms + ev_ms) * 1e-3) / 1e6, "unit": UNIT,
        "fit_us_per_pixel": fit_ms * 1e3 / px, "table_us_per_pixel": ev_ms * 1e3 / px,
        "cg_steps": {"mean": info["steps_mean"], "limit": max_iters, "tol": tol,
                     "pixels_refitted_and_evaluated_in_fp64": unconverged,
                     "why": "not converged, or |w|_2 > 10 tol / 1.2e-7 (float32 distances no longer hold the tolerance)"},
        "roofline": {"bound": "sfu", "kernel": "rbf_eval_f32_kernel",
                     "achieved": G * G * N_LIGHTS * px / (ev_ms * 1e-3) / 1e12, "peak": sm_peak / 1e12,
                     "unit": "Tsqrt/s", "frac": G * G * N_LIGHTS * px / (ev_ms * 1e-3) / sm_peak,
                     "note": "G*G*N square roots per pixel on MUFU; peak = the SFU rate this repo's own microbenchmark "
                             "sustains (tools/peak_sfu.cu: 4.62 T/s = 16 per clock per SM; 4.14 T/s with this "
                             "kernel's instruction mix); MEASURED_PEAKS.json has no SFU figure"},
        "parity_sample": {"pixels": int(len(sample_px)), "grid_max_rel": worst, "tolerance": 1e-3,
                          "table_entries_off_beyond_tolerance": flips},
        "singular_tiles": singular_tiles, "unrepresentable": int(bad[0].item())}
    del tab, state
    # ---- the same on the fp64 path (direct DMMA Cholesky + fp64 evaluation), rows [0, 2)
    px2 = 2 * COLS
    batch = 148 * 2
    nbytes = lib.rti_rbf_workspace_bytes(N_LIGHTS, batch)
    work = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    tab = torch.empty((G, G, px2), dtype=torch.int32, device="cuda")
    bad.zero_()
    w = torch.empty((px2, N_LIGHTS), dtype=torch.float64, device="cuda")
    fit2 = best(lambda: _capi.call("rti_rbf_fit_soa", lx.data_ptr(), ly.data_ptr(), inten.data_ptr(), px2, N_LIGHTS,
                                   w.data_ptr(), work.data_ptr(), nbytes, batch, bad.data_ptr(), S))
    ev2 = best(lambda: _capi.call("rti_rbf_table", lx.data_ptr(), ly.data_ptr(), w.data_ptr(), px2, N_LIGHTS,
                                  d_axis.data_ptr(), G, tab.data_ptr(), px2, bad.data_ptr() + 4, S))
    res["rbf_per_pixel_lights"]["fp64_path"] = {
        "sample": "%d pixels (2 ROI rows)" % px2, "fit_ms": fit2, "table_ms": ev2,
        "value": px2 * N_LIGHTS / ((fit2 + ev2) * 1e-3) / 1e6, "unit": UNIT,
        "fit_fp64_tflops": (N_LIGHTS ** 3 / 3.0) * 2 * px2 / (fit2 * 1e-3) / 1e12,
        "bound": "fp64 tensor pipe (fit; 37 TFLOP/s DMMA peak is this repo's own microbenchmark, "
                 "tools/peak_fp64.cu), fp64 pipe (evaluation)"}
    del lx, ly, inten, w, work, tab
    bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    # ---- shared lights: rows [0, 256) of the slab, nodes = the light vectors seen from the ROI centre
    rows = 256
    px = rows * COLS
    cams = d_cams.cpu().numpy()
    c = np.array([COLS / 2.0, ROWS / 2.0, 0.0])
    l = (cams - c) / np.linalg.norm(cams - c, axis=1, keepdims=True)
    d_lx = torch.from_numpy(l[:, 0].astype(np.float32)).cuda()
    d_ly = torch.from_numpy(l[:, 1].astype(np.float32)).cuda()
    op_rows = G * G
    elems = lib.rti_rbf_shared_operator_elems(op_rows, N_LIGHTS)
    hi = torch.empty(elems, dtype=torch.int16, device="cuda")
    lo = torch.empty(elems, dtype=torch.int16, device="cuda")
    inv = torch.empty(op_rows, dtype=torch.float32, device="cuda")
    packed = torch.empty(lib.rti_rbf_shared_packed_elems(N_LIGHTS, px), dtype=torch.int16, device="cuda")
    tab = torch.empty((op_rows, px), dtype=torch.int32, device="cuda")
    bad.zero_()
    op_ms = best(lambda: _capi.call("rti_rbf_shared_operator", d_lx.data_ptr(), d_ly.data_ptr(), N_LIGHTS,
                                    d_axis.data_ptr(), G, 0, hi.data_ptr(), lo.data_ptr(), inv.data_ptr(),
                                    bad.data_ptr(), S), reps=1)
    pack_ms = best(lambda: _capi.call("rti_rbf_shared_pack_frames", d_frames.data_ptr(), p, N_LIGHTS, px,
                                      packed.data_ptr(), S))
    gemm_ms = best(lambda: _capi.call("rti_rbf_shared_table", packed.data_ptr(), N_LIGHTS, px, hi.data_ptr(),
                                      lo.data_ptr(), inv.data_ptr(), op_rows, tab.data_ptr(), px, bad.data_ptr(), S))
    n_pad, r_pad = -(-N_LIGHTS // 64) * 64, -(-op_rows // 128) * 128
    executed = 4.0 * r_pad * n_pad * px / (gemm_ms * 1e-3) / 1e12
    tensor_peak = None
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            tensor_peak = float(json.load(f)["bf16_tflops"])
    except Exception:
        pass
    res["rbf_shared_lights"] = {
        "sample": "%d pixels (%d ROI rows) x %d lights, pack + tcgen05 contraction into the 100x100 table"
                  % (px, rows, N_LIGHTS),
        "operator_build_ms": op_ms, "pack_ms": pack_ms, "contraction_ms": gemm_ms,
        "value": px * N_LIGHTS / ((pack_ms + gemm_ms) * 1e-3) / 1e6, "unit": UNIT,
        "roofline": {"bound": "tensor", "achieved": executed, "peak": tensor_peak, "unit": "TFLOP/s",
                     "frac": (executed / tensor_peak) if tensor_peak else None,
                     "note": "executed fp16 flops: hi + lo operator halves on the padded shape; peak = measured "
                             "burst bf16 (MEASURED_PEAKS.json bf16_tflops)"},
        "singular": int(bad.item())}
    return res


# ------------------------------------------- BASELINE configs[4]: relighting frames per second
def relight_config5(torch, lib, _capi, hbm_peak):
    """4096 x 4096 PTM coefficient map re-rendered for a sweep of 1000 light directions (the headless
    relighting_event, interactive_relighting.py:11-39): frames/s on one GPU from the fp64 map (bit-identical to
    the table at grid nodes) and the fp32 map, V channel only and with the fused OpenCV HSV->BGR."""
    R = 4096
    p = R * R
    S = torch.cuda.current_stream().cuda_stream
    gen = torch.Generator(device="cuda").manual_seed(3)
    c64 = torch.randn((6, p), dtype=torch.float64, device="cuda", generator=gen) * 40.0
    c32 = torch.empty((6, p), dtype=torch.float32, device="cuda")
    _capi.call("rti_coef_to_f32", c64.data_ptr(), 6 * p, c32.data_ptr(), S)
    hsv = torch.randint(0, 180, (R, R, 3), dtype=torch.uint8, device="cuda", generator=gen)
    dirs = np.random.default_rng(1).uniform(-1, 1, (1000, 2))
    out = {"workload": "4096x4096 PTM coefficient map, 1000 light directions", "frames": len(dirs)}
    for name, fn, cmap, cbytes in (("f64", "rti_relight_ptm", c64, 48), ("f32", "rti_relight_ptm_f32", c32, 24)):
        for label, hsv_t, obytes in (("V", None, 1), ("bgr", hsv, 6)):
            o = torch.empty((R, R, 3) if hsv_t is not None else (R, R), dtype=torch.uint8, device="cuda")
            hp = hsv_t.data_ptr() if hsv_t is not None else None

            def sweep():
                for lu, lv in dirs:
                    _capi.call(fn, cmap.data_ptr(), p, p, R, float(lu), float(lv), hp, o.data_ptr(), S)

            for lu, lv in dirs[:20]:
                _capi.call(fn, cmap.data_ptr(), p, p, R, float(lu), float(lv), hp, o.data_ptr(), S)
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            sweep()
            b.record()
            torch.cuda.synchronize()
            ms = a.elapsed_time(b) / len(dirs)
            gbs = (cbytes + obytes) * p / (ms * 1e-3) / 1e9
            out["%s_%s" % (name, label)] = {"frames_per_s": 1e3 / ms, "ms_per_frame": ms, "achieved_gbs": gbs,
                                            "frac_of_hbm": gbs / hbm_peak}
    out["note"] = ("every frame re-reads the whole map (403 / 805 MB > L2); bound = HBM read; a real-time display "
                   "needs 60 frames/s")
    return out


# ------------------------------------------------------------------ parity sample (checker)
def parity_sample(frames_of, cams, row0, cols, n_pixels, coef_cols, table_cols, idx):
    """Holds sampled pixels of a finished run to the CPU oracle (the checker, after the timed region):
    coefficient columns (6, k) and int32 table columns (G*G, k) of the pixels `idx` (indices into the
    slab) against orc.gather_soa -> ptm_fit -> ptm_grid -> table_from_grids on the same frames.
    frames_of(y, x) returns the (N,) uint8 intensities of slab pixel (y, x)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import rti_oracle as orc
    ax = orc.light_axis()
    max_rel, mismatch, near = 0.0, 0, 0
    for j, pix in enumerate(idx):
        y, x = divmod(int(pix), cols)
        fr = np.asarray(frames_of(y, x), dtype=np.uint8).reshape(-1, 1, 1)
        lx, ly, inten = orc.gather_soa(fr, cams, row0=row0 + y, col0=x)
        ref = orc.ptm_fit(lx[0, 0], ly[0, 0], inten[0, 0])
        if coef_cols is not None:
            max_rel = max(max_rel, float(np.abs(coef_cols[:, j] - ref).max() / np.abs(ref).max()))
        if table_cols is not None:
            grid = orc.ptm_grid(ref, ax)
            tab = orc.table_from_grids(grid[None, None])[:, :, 0, 0].reshape(-1)
            diff = tab != table_cols[:, j]
            # an entry may differ only where the fp64 value sits within 1e-9 of an integer
            frac = np.abs(grid.reshape(-1) - np.rint(grid.reshape(-1)))
            near += int(np.count_nonzero(diff & (frac < 1e-9)))
            mismatch += int(np.count_nonzero(diff & ~(frac < 1e-9)))
    return {"pixels": int(len(idx)), "max_rel": max_rel, "table_mismatch": mismatch,
            "table_integer_boundary_flips": near, "oracle": "oracle/rti_oracle.py (fp64 SVD fit)"}


# ------------------------------------------- BASELINE configs[3], strong-scaled, with the gather
def config4_strong(torch, dist, _capi, lib, rank, world, hbm_peak):
    """synthetic 4096x4096 ROI x 1000 lights, PTM: the ROI rows are split over the ranks
    (sharding.row_partition); every rank streams its slab through the fused kernel in sub-slabs of
    256 rows into ONE reused 42 GB table buffer (the 671 GB table cannot be resident anywhere), then
    the fp64 coefficient map (805 MB in total) is gathered to rank 0 over NCCL (sharding.gather_planar).
    Device-timed, max over ranks; strong scaling (total work fixed)."""
    from rti_b200 import sharding, synthetic
    import rti_b200
    R4, N4, SUB = 4096, 1000, 256
    r0, nr = sharding.row_partition(R4, world)[rank]
    P_loc = nr * R4
    S = torch.cuda.current_stream().cuda_stream
    gen = torch.Generator(device="cuda").manual_seed(100 + rank)
    frames = torch.empty((N4, nr, R4), dtype=torch.uint8, device="cuda")
    for i in range(0, N4, 50):
        frames[i:i + 50] = torch.randint(0, 256, (min(50, N4 - i), nr, R4), dtype=torch.uint8, device="cuda",
                                         generator=gen)
    cams = synthetic.camera_positions(R4, R4, N4, np.random.default_rng(0))
    d_cams = torch.from_numpy(cams).cuda()
    d_axis = torch.from_numpy(rti_b200.light_axis(0.02)).cuda()
    coef = torch.empty((6, P_loc), dtype=torch.float64, device="cuda")
    sub_px = SUB * R4
    table = torch.empty((G * G, sub_px), dtype=torch.int32, device="cuda")
    bad = torch.zeros(2, dtype=torch.int32, device="cuda")

    def run_ptm():
        for y in range(0, nr, SUB):
            rr = min(SUB, nr - y)
            _capi.call("rti_ptm_fit_table_frames", frames.data_ptr() + y * R4, P_loc, d_cams.data_ptr(), N4,
                       r0 + y, rr, R4, d_axis.data_ptr(), G, coef.data_ptr() + 8 * y * R4, P_loc,
                       table.data_ptr(), sub_px, bad.data_ptr(), S)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    run_ptm()                                                        # warm-up
    full = sharding.gather_planar(coef, R4, R4, dst=0) if dist is not None else coef
    barrier()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    ev[0].record()
    run_ptm()
    ev[1].record()
    full = sharding.gather_planar(coef, R4, R4, dst=0) if dist is not None else coef
    ev[2].record()
    barrier()
    t = torch.tensor([ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2]), ev[0].elapsed_time(ev[2])],
                     dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ptm_ms, gather_ms, total_ms = [float(v) for v in t.tolist()]
    # the gather alone, all ranks released together (gather_ms above also contains the wait for the slowest rank)
    ga = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    barrier()
    ga[0].record()
    full = sharding.gather_planar(coef, R4, R4, dst=0) if dist is not None else coef
    ga[1].record()
    barrier()
    t = torch.tensor([ga[0].elapsed_time(ga[1])], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    gather_alone_ms = float(t.item())
    nbad = int(bad.sum().item())
    # sampled oracle check: coefficients anywhere in the slab, table columns in the last sub-slab
    rng = np.random.default_rng(7 + rank)
    last0 = ((nr - 1) // SUB) * SUB
    last_rows = nr - last0
    idx_c = rng.integers(0, P_loc, 6)
    idx_t = rng.integers(0, last_rows * R4, 6)
    fl = frames.view(N4, P_loc)
    par_c = parity_sample(lambda y, x: fl[:, y * R4 + x].cpu().numpy(), cams, r0, R4, P_loc,
                          coef[:, torch.from_numpy(idx_c).cuda()].cpu().numpy(), None, idx_c)
    par_t = parity_sample(lambda y, x: fl[:, (last0 + y) * R4 + x].cpu().numpy(), cams, r0 + last0, R4,
                          last_rows * R4, None, table[:, torch.from_numpy(idx_t).cuda()].cpu().numpy(), idx_t)
    par = torch.tensor([par_c["max_rel"], float(par_t["table_mismatch"]), float(nbad)], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(par, op=dist.ReduceOp.MAX)
    if rank == 0 and dist is not None:
        # the gathered map must hold rank 0's own slab bit for bit
        assert torch.equal(full[:, r0:r0 + nr, :].reshape(6, -1), coef)
    P = R4 * R4
    alg = (N4 + 48 + 4 * G * G) * P
    gather_bytes = 48 * (P - P_loc) if world > 1 else 0

    def max_over_ranks(vals):
        tt = torch.tensor(vals, dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return [float(v) for v in tt.tolist()]

    def timed(fn):
        fn()
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        barrier()
        return a.elapsed_time(b)

    # ---- linear RBF at the same split.  (1) shared lights (nodes = light vectors seen from the ROI centre):
    # operator per rank, then pack + tcgen05 contraction of every sub-slab of the rank's rows (complete run)
    c = np.array([R4 / 2.0, R4 / 2.0, 0.0])
    lv = (cams - c) / np.linalg.norm(cams - c, axis=1, keepdims=True)
    d_lx1 = torch.from_numpy(lv[:, 0].astype(np.float32)).cuda()
    d_ly1 = torch.from_numpy(lv[:, 1].astype(np.float32)).cuda()
    op_rows = G * G
    elems = lib.rti_rbf_shared_operator_elems(op_rows, N4)
    hi = torch.empty(elems, dtype=torch.int16, device="cuda")
    lo = torch.empty(elems, dtype=torch.int16, device="cuda")
    inv = torch.empty(op_rows, dtype=torch.float32, device="cuda")
    packed = torch.empty(lib.rti_rbf_shared_packed_elems(N4, sub_px), dtype=torch.int16, device="cuda")
    sbad = torch.zeros(2, dtype=torch.int32, device="cuda")
    op_ms = timed(lambda: _capi.call("rti_rbf_shared_operator", d_lx1.data_ptr(), d_ly1.data_ptr(), N4, d_axis.data_ptr(),
                                     G, 0, hi.data_ptr(), lo.data_ptr(), inv.data_ptr(), sbad.data_ptr(), S))

    def run_shared():
        for y in range(0, nr, SUB):
            rr = min(SUB, nr - y)
            _capi.call("rti_rbf_shared_pack_frames", frames.data_ptr() + y * R4, P_loc, N4, rr * R4, packed.data_ptr(), S)
            _capi.call("rti_rbf_shared_table", packed.data_ptr(), N4, rr * R4, hi.data_ptr(), lo.data_ptr(),
                       inv.data_ptr(), op_rows, table.data_ptr(), sub_px, sbad.data_ptr(), S)

    shared_ms = timed(run_shared)
    del packed, hi, lo
    # (2) per-pixel lights (the reference's semantics): tile-preconditioned CG fit + float32-distance table on
    # 16 rows of the rank's slab; every pixel is independent and costs the same, so the rank's time is the
    # sample's time scaled by pixels (sampled, extrapolated)
    rs = min(16, nr)
    spx = rs * R4
    lx = torch.empty((spx, N4), dtype=torch.float32, device="cuda")
    ly = torch.empty_like(lx)
    inten = torch.empty((spx, N4), dtype=torch.int32, device="cuda")
    _capi.call("rti_gather_soa", frames.data_ptr(), P_loc, d_cams.data_ptr(), N4, r0, rs, R4,
               lx.data_ptr(), ly.data_ptr(), inten.data_ptr(), S)
    from rti_b200 import analysis
    st4 = {}

    def fit4():
        hb = torch.zeros(1, dtype=torch.int32, device="cuda")
        st4["w"], st4["info"] = analysis.rbf_fit_device_pcg(lx, ly, inten, rs, R4, hb, max_iters=48, tol=0.01)

    tview = table.view(-1)[:G * G * spx].view(G * G, spx)
    fit_ms = timed(fit4)
    w, info4 = st4["w"], st4["info"]
    unconv = info4["refit"]
    ev_ms = timed(lambda: analysis.rbf_table_device(lx, ly, w, d_axis, G, tview, sbad, fp64_idx=info4["fp64_idx"]))
    # one sampled pixel per rank against scipy's algorithm in fp64
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import rti_oracle as orc
    q = int(rng.integers(0, spx))
    l1, l2, dd = lx[q].cpu().numpy(), ly[q].cpu().numpy(), inten[q].cpu().numpy()
    gref = orc.rbf_grid(l1, l2, orc.rbf_fit(l1, l2, dd), orc.light_axis())
    ggot = orc.rbf_grid(l1, l2, w[q].cpu().numpy(), orc.light_axis())
    rbf_rel = float(np.max(np.abs(ggot - gref)) / np.max(np.abs(gref)))
    steps_mean = info4["steps_mean"]
    op_ms, shared_ms, fit_ms, ev_ms, rbf_rel, unconv_max = max_over_ranks([op_ms, shared_ms, fit_ms, ev_ms, rbf_rel, unconv])
    pp_s = (fit_ms + ev_ms) * (P_loc / spx) / 1e3
    n_pad, r_pad = -(-N4 // 64) * 64, -(-op_rows // 128) * 128
    rbf = {
        "shared_lights": {"ms": shared_ms, "operator_build_ms": op_ms, "value": P * N4 / (shared_ms * 1e-3) / 1e6,
                          "unit": UNIT, "executed_tflops_per_gpu": 4.0 * r_pad * n_pad * P_loc / (shared_ms * 1e-3) / 1e12,
                          "what": "pack + tcgen05 contraction into the int32 table, complete run of the rank's rows"},
        "per_pixel_lights": {"sample_pixels_per_rank": spx, "fit_ms": fit_ms, "table_ms": ev_ms,
                             "cg_steps_mean": steps_mean, "pixels_in_fp64": int(unconv_max),
                             "extrapolated_full_roi_s": pp_s, "value": P * N4 / pp_s / 1e6, "unit": UNIT,
                             "grid_max_rel_vs_oracle": rbf_rel,
                             "what": "pcg fit + float32-distance table (fp64 kernels for the pixels the fit reports) on a "
                                     "16-row sample per rank; sampled, extrapolated"}}
    del lx, ly, inten, w
    out = {"workload": "synthetic 4096x4096 ROI x 1000 light samples, PTM fit + 100x100 light grid (BASELINE configs[3])",
           "scaling": "strong", "ranks": world, "rows_per_rank": nr, "sub_slab_rows": SUB,
           "ptm_ms": ptm_ms, "gather_ms": gather_ms, "total_ms": total_ms,
           "value": P * N4 / (total_ms * 1e-3) / 1e6, "unit": UNIT,
           "nccl_gather_bytes": gather_bytes,
           "gather_alone_ms": gather_alone_ms,
           "nccl_gather_gbs": (gather_bytes / (gather_alone_ms * 1e-3) / 1e9) if world > 1 else None,
           "gather": "fp64 coefficient map (6, rows, cols) to rank 0, sharding.gather_planar (batch_isend_irecv)",
           "roofline": {"bound": "hbm", "achieved": alg / world / (ptm_ms * 1e-3) / 1e9, "peak": hbm_peak,
                        "frac": alg / world / (ptm_ms * 1e-3) / 1e9 / hbm_peak, "unit": "GB/s per GPU"},
           "parity_sample": {"pixels_per_rank": 12, "max_rel": float(par[0].item()),
                             "table_mismatch": int(par[1].item()), "singular": int(par[2].item())},
           "rbf": rbf}
    del frames, coef, table, full
    torch.cuda.empty_cache()
    return out


# ------------------------------------------------------------------------------ GPU arm
def run_gpu(args, rank, local_rank, world):
    import torch
    import rti_b200
    from rti_b200 import _capi, synthetic

    cpu = cpu_rbf = None
    if world == 1 and not args.no_cpu_baseline:       # before CUDA is initialised: the pool forks
        cpu, _ = cpu_reference(args.cpu_pixels_per_core)
        cpu_rbf, _ = cpu_reference(max(args.cpu_pixels_per_core // 12, 1), mode="rbf")

    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    lib = _capi.load()
    rows_total = ROWS * world
    row0 = ROWS * rank
    p = ROWS * COLS
    units_per_rank = p * N_LIGHTS

    # synthetic inputs: uniform-random uint8 frames of this rank's slab, shared camera track
    frames, cams = synthetic.random_stack(ROWS, COLS, N_LIGHTS, seed=1 + rank, rows_total=rows_total)
    cams = synthetic.camera_positions(rows_total, COLS, N_LIGHTS, np.random.default_rng(0))
    axis = rti_b200.light_axis(0.02)
    h_frames = torch.from_numpy(frames).pin_memory()
    d_frames = h_frames.cuda()
    d_cams = torch.from_numpy(cams).cuda()
    d_axis = torch.from_numpy(axis).cuda()
    d_coef = torch.empty((6, p), dtype=torch.float64, device="cuda")
    d_table = torch.empty((G, G, p), dtype=torch.int32, device="cuda")          # 42 GB
    d_bad = torch.zeros(2, dtype=torch.int32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream

    def step():
        # one fused kernel: gather + fit + table (the two-kernel form is timed below for comparison)
        _capi.call("rti_ptm_fit_table_frames", d_frames.data_ptr(), p, d_cams.data_ptr(), N_LIGHTS, row0, ROWS, COLS,
                   d_axis.data_ptr(), G, d_coef.data_ptr(), p, d_table.data_ptr(), p, d_bad.data_ptr(), stream)

    def step_two_kernels():
        _capi.call("rti_ptm_fit_frames", d_frames.data_ptr(), p, d_cams.data_ptr(), N_LIGHTS, row0, ROWS, COLS,
                   d_coef.data_ptr(), p, d_bad.data_ptr(), stream)
        _capi.call("rti_ptm_table", d_coef.data_ptr(), p, p, d_axis.data_ptr(), G, d_table.data_ptr(), p,
                   d_bad.data_ptr() + 4, stream)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    assert d_bad.sum().item() == 0, "singular / unrepresentable pixels in the synthetic workload"
    d_bad.zero_()

    # two-kernel form (untimed region, for the comparison printed under "kernel_ms")
    ev2 = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    step_two_kernels()
    ev2[0].record()
    _capi.call("rti_ptm_fit_frames", d_frames.data_ptr(), p, d_cams.data_ptr(), N_LIGHTS, row0, ROWS, COLS,
               d_coef.data_ptr(), p, d_bad.data_ptr(), stream)
    ev2[1].record()
    _capi.call("rti_ptm_table", d_coef.data_ptr(), p, p, d_axis.data_ptr(), G, d_table.data_ptr(), p,
               d_bad.data_ptr() + 4, stream)
    ev2[2].record()
    barrier()
    fit_ms, table_ms = ev2[0].elapsed_time(ev2[1]), ev2[1].elapsed_time(ev2[2])

    # ---- timed region: exactly K steps, device-timed, per-launch events for the roofline
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(args.steps)]
    sampler = ClockSampler(local_rank)
    lib.rti_launch_count(1)
    sampler.start()
    barrier()
    t_start = torch.cuda.Event(enable_timing=True)
    t_end = torch.cuda.Event(enable_timing=True)
    t_start.record()
    for k in range(args.steps):
        ev[k][0].record()
        step()
        ev[k][1].record()
    t_end.record()
    barrier()
    clocks = sampler.stop()
    launches = int(lib.rti_launch_count(0))
    total_ms = t_start.elapsed_time(t_end)
    fused_ms = float(np.mean([e[0].elapsed_time(e[1]) for e in ev]))
    if dist is not None:
        t = torch.tensor([total_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
    ms_per_step = total_ms / args.steps
    value = units_per_rank * world / (ms_per_step * 1e-3) / 1e6

    # ---- sampled oracle check of what the timed steps left in HBM (coefficients + table columns)
    rng = np.random.default_rng(11 + rank)
    idx = rng.integers(0, p, args.parity_pixels)
    d_idx = torch.from_numpy(idx).cuda()
    parity = parity_sample(lambda y, x: frames[:, y, x], cams, row0, COLS, p,
                           d_coef[:, d_idx].cpu().numpy(), d_table.view(G * G, p)[:, d_idx].cpu().numpy(), idx)
    parity["singular"] = int(d_bad.sum().item())

    # ---- e2e: the host-buffer C-ABI call, H2D of the frames and D2H of the table inside
    e2e = None
    table_bytes = G * G * p * 4
    if not args.no_e2e:
        d_table = None
        torch.cuda.empty_cache()
        try:
            import psutil
            avail = psutil.virtual_memory().available
        except Exception:
            avail = 1 << 62
        if avail < world * table_bytes * 1.3:
            e2e = {"value": None, "unit": UNIT, "skipped": "host RAM: %.0f GB available, need %.0f GB pinned"
                   % (avail / 1e9, world * table_bytes / 1e9)}
        else:
            h_table = torch.empty((G, G, p), dtype=torch.int32, pin_memory=True)
            h_cams = torch.from_numpy(cams)

            def host_step():
                _capi.call("rti_ptm_compute_host", h_frames.data_ptr(), h_cams.data_ptr(), N_LIGHTS, ROWS, COLS,
                           0, ROWS, axis.ctypes.data, G, h_table.data_ptr(), None, 0)

            # the slab's global row offset only shifts y; rti_ptm_compute_host takes row0 within the
            # host frame buffer, so ranks pass their slab as a whole ROI of ROWS rows here
            host_step()                                                  # warm-up (allocates the workspace)
            barrier()
            t0 = time.perf_counter()
            for _ in range(args.e2e_steps):
                host_step()
            barrier()
            dt = (time.perf_counter() - t0) / args.e2e_steps
            if dist is not None:
                t = torch.tensor([dt], dtype=torch.float64, device="cuda")
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                dt = float(t.item())
            # bytes that crossed PCIe in the last call: the table's 4 B per entry, or 2 B when the library moved
            # it as int16 and widened it on the host (rti_b200.h: narrow transfer; same int32 table either way)
            d2h_bytes = int(lib.rti_host_d2h_bytes())
            e2e = {"value": units_per_rank * world / dt / 1e6, "unit": UNIT,
                   "h2d_bytes_per_step": int(frames.nbytes + cams.nbytes + axis.nbytes),
                   "d2h_bytes_per_step": d2h_bytes, "ms_per_step": dt * 1e3, "steps": args.e2e_steps,
                   "table_bytes_delivered_per_step": int(table_bytes),
                   "transfer": "int32 over PCIe" if d2h_bytes >= table_bytes else
                   "int16 over PCIe, widened to int32 by host threads" if d2h_bytes <= table_bytes // 2 else
                   "%.0f %% of the table as int16 over PCIe, widened to int32 by host threads; the rest as int32"
                   % (100.0 * (table_bytes - d2h_bytes) / (table_bytes / 2)),
                   "api": "rti_ptm_compute_host (pinned host frames -> int32 table in pinned host memory)"}
            lib.rti_destroy()
            # the host table of the last e2e step against the oracle (the host call numbers rows from 0)
            hp = parity_sample(lambda y, x: frames[:, y, x], cams, 0, COLS, p, None,
                               h_table.view(G * G, p)[:, torch.from_numpy(idx)].numpy(), idx)
            e2e["parity_sample"] = {"pixels": hp["pixels"], "table_mismatch": hp["table_mismatch"],
                                    "table_integer_boundary_flips": hp["table_integer_boundary_flips"]}
            # ---- e2e roofline: the same D2H bytes as bare pinned copies (4 GiB pieces of a device
            # buffer into consecutive parts of the pinned table), all ranks at the same time
            piece = min(1 << 32, table_bytes)
            d_src = torch.empty(piece, dtype=torch.uint8, device="cuda")
            h_flat = h_table.view(torch.uint8).view(-1)
            barrier()
            t0 = time.perf_counter()
            off = 0
            while off < table_bytes:
                nb = min(piece, table_bytes - off)
                h_flat[off:off + nb].copy_(d_src[:nb], non_blocking=True)
                off += nb
            torch.cuda.synchronize()
            bare = time.perf_counter() - t0
            if dist is not None:
                t = torch.tensor([bare], dtype=torch.float64, device="cuda")
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                bare = float(t.item())
            e2e["roofline"] = {"bound": "pcie d2h", "what": "bare cudaMemcpyAsync of the step's D2H bytes into the same "
                               "pinned buffer, %d rank(s) concurrently" % world,
                               "bare_copy_ms": bare * 1e3, "bare_copy_gbs_per_rank": table_bytes / bare / 1e9,
                               "bare_copy_gbs_aggregate": world * table_bytes / bare / 1e9,
                               "e2e_gbs_per_rank": d2h_bytes / dt / 1e9,
                               "table_gbs_delivered_per_rank": table_bytes / dt / 1e9,
                               "frac": (d2h_bytes / dt) / (table_bytes / bare),
                               "note": "frac = PCIe bytes actually moved per second / the bare copy rate; with the "
                               "int16 transfer the rest of the step is the host threads' int32 stores"}
            del d_src, h_table

    secondary = None
    if world == 1 and not args.no_secondary:
        secondary = secondary_rbf(torch, lib, _capi, d_frames, d_cams, d_axis, p)
        if cpu_rbf is not None:
            secondary["rbf_per_pixel_lights"]["cpu_baseline"] = cpu_rbf
            secondary["rbf_per_pixel_lights"]["vs_cpu"] = secondary["rbf_per_pixel_lights"]["value"] / cpu_rbf["value"]
    relight = None
    if world == 1 and not args.no_secondary:
        d_table = d_coef = d_frames = None
        torch.cuda.empty_cache()
        relight = relight_config5(torch, lib, _capi, read_peaks()[0])
    config4 = None
    if not args.no_config4:
        d_table = d_coef = d_frames = None                          # release the headline buffers
        torch.cuda.empty_cache()
        config4 = config4_strong(torch, dist, _capi, lib, rank, world, read_peaks()[0])

    if rank != 0:
        return

    # ---- roofline of the step's kernel (gather + fit + table fused): HBM-bound on the table write
    peak, peak_src = read_peaks()
    alg_bytes = table_bytes + units_per_rank + 48 * p        # int32 table + uint8 frames + coefficient map
    achieved = alg_bytes / (fused_ms * 1e-3) / 1e9
    tbl_bytes = table_bytes + 48 * p
    roofline = {"bound": "hbm", "kernel": "ptm_fused_ws_kernel<true>", "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "traffic": read_traffic(), "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg_bytes, "launch_ms": fused_ms,
                "two_kernel_form": {
                    "ptm_fit_frames_kernel": {"launch_ms": fit_ms, "bound": "fp64 pipe",
                                              "units_per_launch": units_per_rank,
                                              "gunits_per_s": units_per_rank / (fit_ms * 1e-3) / 1e9},
                    "ptm_table_kernel<4>": {"launch_ms": table_ms, "bound": "hbm",
                                            "achieved": tbl_bytes / (table_ms * 1e-3) / 1e9,
                                            "frac": tbl_bytes / (table_ms * 1e-3) / 1e9 / peak}}}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": make_config(world),
        "e2e": e2e, "parity_sample": parity, "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
        "kernel_ms": {"ptm_fused": fused_ms, "ptm_fit_frames": fit_ms, "ptm_table": table_ms},
        "secondary": secondary, "config5_relight": relight, "config4": config4,
    }
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="skip the linear-RBF side measurements")
    ap.add_argument("--no-config4", action="store_true", help="skip the strong-scaled BASELINE configs[3] run")
    ap.add_argument("--parity-pixels", type=int, default=64)
    ap.add_argument("--cpu-pixels-per-core", type=int, default=96)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "native" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            # launched without torchrun: re-launch one process per GPU
            import subprocess
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus),
                   "--master-addr", "127.0.0.1", "--master-port", "29531", os.path.abspath(__file__)] + sys.argv[1:]
            sys.exit(subprocess.call(cmd))
        raise SystemExit("WORLD_SIZE=%d but --gpus %d" % (world, args.gpus))
    # Native libraries (NCCL prints its version banner) write to fd 1; the contract is ONE JSON line on
    # stdout, so everything but the final line goes to stderr.
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    try:
        line = run_gpu(args, rank, local_rank, world)
    finally:
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        os.close(real_stdout)
    if line is not None:
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
