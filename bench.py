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
def _cpu_worker(args):
    """Oracle port of the per-pixel path for `count` pixels: full-SVD PTM fit
    (analysis.py:295-298), grid evaluation (:301-315) and int32 truncation (:404-407)."""
    seed, count, n = args
    os.environ["OMP_NUM_THREADS"] = "1"
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import rti_oracle as orc
    rng = np.random.default_rng(seed)
    lx = rng.uniform(-0.9, 0.9, (count, n)).astype(np.float32)
    ly = rng.uniform(-0.9, 0.9, (count, n)).astype(np.float32)
    inten = rng.integers(0, 256, (count, n)).astype(np.int32)
    ax = orc.light_axis()
    t0 = time.perf_counter()
    acc = 0
    for p in range(count):
        grid = orc.ptm_grid(orc.ptm_fit(lx[p], ly[p], inten[p]), ax)
        acc += int(orc.table_from_grids(grid[None, None])[0, 0, 0, 0])
    return time.perf_counter() - t0, acc


def cpu_reference(pixels_per_core, steps=1, warmup=0):
    """Times the oracle port on all host cores; returns (Mpixel.lights/s, cores, sample text, ms/step)."""
    import multiprocessing as mp
    try:
        import threadpoolctl
        threadpoolctl.threadpool_limits(1)
    except Exception:
        pass
    cores = len(os.sched_getaffinity(0))
    ctx = mp.get_context("fork")
    with ctx.Pool(cores) as pool:
        pool.map(_cpu_worker, [(i, 1, N_LIGHTS) for i in range(cores)])          # warm the workers
        for _ in range(warmup):
            pool.map(_cpu_worker, [(i, pixels_per_core, N_LIGHTS) for i in range(cores)])
        t0 = time.perf_counter()
        for s in range(max(steps, 1)):
            pool.map(_cpu_worker, [(1000 * s + i, pixels_per_core, N_LIGHTS) for i in range(cores)])
        dt = (time.perf_counter() - t0) / max(steps, 1)
    pixels = pixels_per_core * cores
    value = pixels * N_LIGHTS / dt / 1e6
    sample = ("%d random pixels x %d lights per step (%d per core), oracle port of full-SVD PTM fit + "
              "100x100 grid eval + int32 truncation; extrapolates linearly in pixels"
              % (pixels, N_LIGHTS, pixels_per_core))
    return value, cores, sample, dt * 1e3


def run_reference(args, rank):
    if rank != 0:
        return
    value, cores, sample, ms = cpu_reference(args.cpu_pixels_per_core, steps=args.steps, warmup=min(args.warmup, 1))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "rows": ROWS, "cols": COLS, "lights": N_LIGHTS, "grid": [G, G]},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
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
    bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    # ---- per-pixel lights: rows [0, 2) of the slab
    px = 2 * COLS
    lx = torch.empty((px, N_LIGHTS), dtype=torch.float32, device="cuda")
    ly = torch.empty_like(lx)
    inten = torch.empty((px, N_LIGHTS), dtype=torch.int32, device="cuda")
    _capi.call("rti_gather_soa", d_frames.data_ptr(), p, d_cams.data_ptr(), N_LIGHTS, 0, 2, COLS,
               lx.data_ptr(), ly.data_ptr(), inten.data_ptr(), S)
    w = torch.empty((px, N_LIGHTS), dtype=torch.float64, device="cuda")
    batch = 148 * 2
    nbytes = lib.rti_rbf_workspace_bytes(N_LIGHTS, batch)
    work = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    tab = torch.empty((G, G, px), dtype=torch.int32, device="cuda")
    fit_ms = best(lambda: _capi.call("rti_rbf_fit_soa", lx.data_ptr(), ly.data_ptr(), inten.data_ptr(), px, N_LIGHTS,
                                     w.data_ptr(), work.data_ptr(), nbytes, batch, bad.data_ptr(), S))
    ev_ms = best(lambda: _capi.call("rti_rbf_table", lx.data_ptr(), ly.data_ptr(), w.data_ptr(), px, N_LIGHTS,
                                    d_axis.data_ptr(), G, tab.data_ptr(), px, bad.data_ptr(), S))
    res["rbf_per_pixel_lights"] = {
        "sample": "%d pixels (2 ROI rows) x %d lights, fit + 100x100 table" % (px, N_LIGHTS),
        "fit_ms": fit_ms, "table_ms": ev_ms, "value": px * N_LIGHTS / ((fit_ms + ev_ms) * 1e-3) / 1e6, "unit": UNIT,
        "fit_fp64_tflops": (N_LIGHTS ** 3 / 3.0) * 2 * px / (fit_ms * 1e-3) / 1e12,
        "bound": "fp64 tensor pipe (fit), fp64 pipe (evaluation)", "singular": int(bad.item())}
    del lx, ly, inten, w, work, tab
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


# ------------------------------------------------------------------------------ GPU arm
def run_gpu(args, rank, local_rank, world):
    import torch
    import rti_b200
    from rti_b200 import _capi, synthetic

    cpu = None
    if world == 1 and not args.no_cpu_baseline:       # before CUDA is initialised: the pool forks
        v, cores, sample, _ = cpu_reference(args.cpu_pixels_per_core)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}

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

    # ---- e2e: the host-buffer C-ABI call, H2D of the frames and D2H of the table inside
    e2e = None
    table_bytes = G * G * p * 4
    if not args.no_e2e:
        del d_table
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
            e2e = {"value": units_per_rank * world / dt / 1e6, "unit": UNIT,
                   "h2d_bytes_per_step": int(frames.nbytes + cams.nbytes + axis.nbytes),
                   "d2h_bytes_per_step": int(table_bytes), "ms_per_step": dt * 1e3, "steps": args.e2e_steps,
                   "api": "rti_ptm_compute_host (pinned host frames -> int32 table in pinned host memory)"}
            lib.rti_destroy()

    secondary = None
    if world == 1 and not args.no_secondary:
        secondary = secondary_rbf(torch, lib, _capi, d_frames, d_cams, d_axis, p)

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
        "config": {"workload": WORKLOAD, "rows_per_gpu": ROWS, "cols": COLS, "lights": N_LIGHTS,
                   "grid": [G, G], "parallelism": "row-tile x%d, no collective" % world,
                   "l2": "inputs (524 MB) and the 42 GB table written per step exceed the 126 MB L2"},
        "e2e": e2e, "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
        "kernel_ms": {"ptm_fused": fused_ms, "ptm_fit_frames": fit_ms, "ptm_table": table_ms},
        "secondary": secondary,
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
