#!/usr/bin/env python
"""BASELINE.json configs[0]/[1] stand-in, end to end through the drop-in call surface:
`compute(video_name, from_storage=True, interpolate_PTM=...)` (analysis.py:414-482) at the reference's
default ROI (400 x 400) on a synthetic `frames_results_<name>.pbz2` with the reference's pickle layout
(the real coin1-4 blobs are not in the reference mount).  Wall-clock per stage:

  read      ut.read_from_file     bz2 + unpickle of the frame list (parallel block decoder)
  step2     compute_intensities   records -> pinned slab -> HBM
  step3     interpolate_intensities   PTM fit / linear-RBF fit on the GPU
  step4+write   prepare_images_data + ut.write_on_file fused: table on the GPU, device -> pinned ring ->
            result pickle, image by image (6.4 GB)
  separate_calls_s: read + compute_intensities + interpolate_intensities + prepare_images_data as
            separate public calls with the table assembled in host memory (what install() gives)

and a check of the written file: `pickle.load` gives list[100] of list[100] of writable int32 (R, R)
arrays equal to the table (what interactive_relighting.py:95 reads).

    python tools/bench_compute.py [--roi 400] [--lights 300] [--json profiles/rNN_compute.json] [--keep]
"""
import argparse
import json
import os
import pickle
import shutil
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import rti_b200  # noqa: E402
from rti_b200 import analysis, synthetic  # noqa: E402


class StageTimer:
    """Wraps module-level functions of rti_b200.analysis and records their wall-clock."""

    def __init__(self):
        self.times = {}

    def wrap(self, module, name, label):
        fn = getattr(module, name)

        def timed(*a, **kw):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            out = fn(*a, **kw)
            torch.cuda.synchronize()
            self.times[label] = self.times.get(label, 0.0) + time.perf_counter() - t0
            return out

        setattr(module, name, timed)
        return fn


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--roi", type=int, default=400)
    ap.add_argument("--lights", type=int, default=300)
    ap.add_argument("--json", default=None)
    ap.add_argument("--dir", default=None, help="working directory (default: a temporary one)")
    ap.add_argument("--skip-rbf", action="store_true")
    ap.add_argument("--big", action="store_true",
                    help="large ROI (e.g. 1024 x 500 lights: 42 GB result): skip the whole-table reloads and the "
                         "separate-calls route, check sampled images through the mmap loader only")
    args = ap.parse_args()
    torch.cuda.set_device(0)
    r, n = args.roi, args.lights
    work = args.dir or tempfile.mkdtemp(prefix="rti_compute_")
    os.makedirs(os.path.join(work, "assets"), exist_ok=True)
    cwd = os.getcwd()
    os.chdir(work)
    out = {"config": "synthetic frames_results_synth.pbz2, ROI %d x %d, %d light samples, 100x100 light grid" % (r, r, n),
           "units_pixel_lights": r * r * n}
    need = r * r * 10000 * 4 * 1.05
    free = shutil.disk_usage(work).free
    if free < need:
        print(json.dumps({"skipped": "result file needs %.1f GB, %.1f GB free under %s" % (need / 1e9, free / 1e9, work)}))
        return
    try:
        t0 = time.perf_counter()
        frames, cams = synthetic.lambertian_stack(r, r, n, seed=0)
        rti_b200.write_on_file(synthetic.as_records(frames, cams), "assets/frames_results_synth", compressed=True)
        out["fixture_s"] = time.perf_counter() - t0
        out["pbz2_bytes"] = os.path.getsize("assets/frames_results_synth.pbz2")
        rti_b200.constants.ROI_DIAMETER = r
        for mode, ptm in (("ptm", True), ("rbf", False)):
            if not ptm and args.skip_rbf:
                continue
            st = StageTimer()
            originals = {k: st.wrap(analysis, k, lab) for k, lab in
                         (("read_from_file", "read"), ("compute_intensities", "step2"),
                          ("interpolate_intensities", "step3"), ("write_table_file", "step4+write"))}
            try:
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                analysis.compute(video_name="synth", from_storage=True, interpolate_PTM=ptm)
                total = time.perf_counter() - t0
            finally:
                for k, fn in originals.items():
                    setattr(analysis, k, fn)
            path = "assets/interpolation_results_synth.pickle"
            size = os.path.getsize(path)
            if args.big:
                t0 = time.perf_counter()
                mapped = analysis.read_from_file("assets/interpolation_results_synth", compressed=False)
                map_s = time.perf_counter() - t0
                data = analysis.compute_intensities(analysis.read_from_file("assets/frames_results_synth"))
                res = analysis.interpolate_intensities(data, interpolate_PTM=ptm)
                ok = len(mapped) == 100 and len(mapped[0]) == 100 and mapped[3][7].shape == (r, r)
                for y in (0, r // 2, r - 1):                        # three ROI rows of four images, re-evaluated
                    gr = res.grids(y * r, r)                        # (r, 100, 100) fp64
                    for v, u in ((0, 0), (3, 7), (50, 50), (99, 99)):
                        want = np.trunc(gr[:, v, u]).astype(np.int32)
                        if ptm:
                            ok = ok and bool(np.array_equal(mapped[v][u][y], want))
                        else:
                            # the default linear-RBF path is float32-distance (north_star: 1e-3): an entry may differ
                            # by one where the fp64 value lies within that tolerance of an integer
                            d = mapped[v][u][y] != want
                            near = np.abs(gr[:, v, u] - np.rint(gr[:, v, u])) <= 1e-3 * np.abs(gr).max()
                            ok = ok and bool(np.all(np.abs(mapped[v][u][y] - want) <= 1) and np.all(near[d]))
                out[mode] = {"total_s": total, "stages_s": st.times, "result_bytes": size, "mapped_load_s": map_s,
                             "result_file_ok": bool(ok), "path": res.path,
                             "mpixel_lights_per_s_wall": r * r * n / total / 1e6}
                print(mode, json.dumps(out[mode]), flush=True)
                del mapped
                os.remove(path)
                continue
            t0 = time.perf_counter()
            with open(path, "rb") as f:
                back = pickle.load(f)
            load_s = time.perf_counter() - t0
            t0 = time.perf_counter()
            mapped = analysis.read_from_file("assets/interpolation_results_synth", compressed=False)   # mmap loader
            mapped_ok = bool(np.array_equal(mapped[50][50], back[50][50]) and mapped[50][50].flags.writeable)
            map_s = time.perf_counter() - t0
            del mapped
            # the same table again through the separate public functions (the install() route)
            t0 = time.perf_counter()
            images = analysis.prepare_images_data(analysis.interpolate_intensities(
                analysis.compute_intensities(analysis.read_from_file("assets/frames_results_synth")), interpolate_PTM=ptm))
            separate_s = time.perf_counter() - t0
            t0 = time.perf_counter()
            analysis.write_on_file(images, "assets/install_route", compressed=False)     # ut.write_on_file after install()
            install_write_s = time.perf_counter() - t0
            same_file = os.path.getsize("assets/install_route.pickle") == size
            os.remove("assets/install_route.pickle")
            ok = (len(back) == 100 and len(back[0]) == 100 and back[3][7].dtype == np.int32
                  and back[3][7].shape == (r, r) and back[3][7].flags.writeable
                  and all(np.array_equal(back[v][u], images[v][u]) for v, u in ((0, 0), (3, 7), (50, 50), (99, 99))))
            out[mode] = {"total_s": total, "stages_s": st.times, "result_bytes": size, "reload_s": load_s, "mapped_load_s": map_s, "mapped_ok": mapped_ok,
                         "separate_calls_s": separate_s, "install_route_write_s": install_write_s,
                         "install_route_same_size": bool(same_file), "result_file_ok": bool(ok),
                         "mpixel_lights_per_s_wall": r * r * n / total / 1e6}
            print(mode, json.dumps(out[mode]), flush=True)
            del back, images
            os.remove(path)
    finally:
        os.chdir(cwd)
        if args.dir is None:
            shutil.rmtree(work, ignore_errors=True)
    if args.json:
        with open(os.path.join(ROOT, args.json), "w") as f:
            json.dump(out, f, indent=1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
