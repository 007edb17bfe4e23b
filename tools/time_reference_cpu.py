#!/usr/bin/env python
"""Times the UNMODIFIED reference functions next to the oracle port on this machine's CPU (one core), for the
per-pixel path bench.py's CPU arm measures: `_interpolate_PTM` / `_interpolate_RBF` (analysis.py:263-317, 249-260)
+ the int32 truncation of prepare_images_data (:404-407), and `compute_intensities` (:196-246) per pixel.light.
The reference is imported from where it is mounted (RTI_REFERENCE_DIR, default /root/reference) and never copied;
the GPU box has no reference, so bench.py times the port there and this file gives the port : reference ratio.

    python tools/time_reference_cpu.py [--n 500] [--pixels 16] [--json profiles/rNN_reference_vs_port_cpu.json]
"""
import argparse
import json
import os
import sys
import time
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
os.environ.setdefault("OMP_NUM_THREADS", "1")
import rti_oracle as orc  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=500)
    ap.add_argument("--pixels", type=int, default=16)
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    ref_dir = os.environ.get("RTI_REFERENCE_DIR", "/root/reference")
    if not os.path.isfile(os.path.join(ref_dir, "analysis.py")):
        print(json.dumps({"unavailable": "reference not mounted at %s" % ref_dir}))
        return
    try:
        import threadpoolctl
        threadpoolctl.threadpool_limits(1)
    except Exception:
        pass
    mpl, plt = types.ModuleType("matplotlib"), types.ModuleType("matplotlib.pyplot")
    mpl.pyplot = plt
    sys.modules.setdefault("matplotlib", mpl)
    sys.modules.setdefault("matplotlib.pyplot", plt)
    sys.path.insert(0, ref_dir)
    cwd = os.getcwd()
    os.chdir(ref_dir)
    try:
        import analysis
        import constants
    finally:
        os.chdir(cwd)
    n, px = args.n, args.pixels
    rng = np.random.default_rng(0)
    lx = rng.uniform(-0.9, 0.9, (px, n)).astype(np.float32)
    ly = rng.uniform(-0.9, 0.9, (px, n)).astype(np.float32)
    inten = rng.integers(0, 256, (px, n)).astype(np.int32)
    ax = orc.light_axis()
    xi, yi = np.meshgrid(ax, ax, indexing="ij")
    out = {"lights": n, "pixels": px, "cores": 1, "python": sys.version.split()[0], "numpy": np.__version__}

    def timed(fn):
        fn(0)
        t0 = time.perf_counter()
        for p in range(px):
            fn(p)
        return (time.perf_counter() - t0) / px

    def trunc(g):
        return int(orc.table_from_grids(np.asarray(g)[None, None])[0, 0, 0, 0])

    t_ref = timed(lambda p: trunc(analysis._interpolate_PTM(lx[p], ly[p], ax, inten[p])))
    t_port = timed(lambda p: trunc(orc.ptm_grid(orc.ptm_fit(lx[p], ly[p], inten[p]), ax)))
    out["ptm"] = {"reference_ms_per_pixel": t_ref * 1e3, "port_ms_per_pixel": t_port * 1e3,
                  "reference_mpixel_lights_per_s": n / t_ref / 1e6, "port_mpixel_lights_per_s": n / t_port / 1e6,
                  "port_over_reference": t_ref / t_port}
    t_ref = timed(lambda p: trunc(analysis._interpolate_RBF(lx[p], ly[p], xi, yi, inten[p])))
    t_port = timed(lambda p: trunc(orc.rbf_grid(lx[p], ly[p], orc.rbf_fit(lx[p], ly[p], inten[p]), ax)))
    out["rbf"] = {"reference_ms_per_pixel": t_ref * 1e3, "port_ms_per_pixel": t_port * 1e3,
                  "reference_mpixel_lights_per_s": n / t_ref / 1e6, "port_mpixel_lights_per_s": n / t_port / 1e6,
                  "port_over_reference": t_ref / t_port}
    # compute_intensities on a small ROI (the triple Python loop, analysis.py:221-232)
    r, nf = 12, 40
    constants.ROI_DIAMETER = r
    recs = [(rng.integers(0, 256, (r, r), dtype=np.uint8), rng.uniform(-r, 2 * r, 3) + np.array([0, 0, 2.0 * r]))
            for _ in range(nf)]
    t0 = time.perf_counter()
    analysis.compute_intensities(recs)
    t_ref = (time.perf_counter() - t0) / (r * r * nf)
    frames = np.stack([f for f, _ in recs])
    cams = np.stack([c for _, c in recs])
    t0 = time.perf_counter()
    orc.gather_soa(frames, cams)
    t_port = (time.perf_counter() - t0) / (r * r * nf)
    out["compute_intensities"] = {"reference_us_per_pixel_light": t_ref * 1e6, "port_us_per_pixel_light": t_port * 1e6}
    print(json.dumps(out))
    if args.json:
        with open(os.path.join(ROOT, args.json), "w") as f:
            json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
