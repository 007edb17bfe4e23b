#!/usr/bin/env python
"""A COMPLETE linear-RBF run (per-pixel lights, the reference's default mode) of one rank's share of BASELINE
configs[3] on one GPU through the drop-in's banded route: rows x 4096 pixels x 1000 lights (default 512 rows = what
each of 8 ranks owns), gather + pcg fit + float32-distance table per band of ROI rows, the table tiles produced into
the two bounded device buffers and dropped (a 4096-wide table cannot be kept anywhere).  Not sampled, not
extrapolated.  Also checks a few pixels' fp64 grids against the oracle.

    python tools/bench_rbf_banded.py [--rows 512] [--cols 4096] [--n 1000] [--json profiles/rNN_rbf_banded.json]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import torch  # noqa: E402

import rti_b200  # noqa: E402
from rti_b200 import analysis, synthetic  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=512)
    ap.add_argument("--cols", type=int, default=4096)
    ap.add_argument("--n", type=int, default=1000)
    ap.add_argument("--roi", type=int, default=4096)
    ap.add_argument("--row0", type=int, default=1536)
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    torch.cuda.set_device(0)
    rows, cols, n = args.rows, args.cols, args.n
    gen = torch.Generator(device="cuda").manual_seed(0)
    frames = torch.randint(0, 256, (n, rows, cols), dtype=torch.uint8, device="cuda", generator=gen)
    cams = synthetic.camera_positions(args.roi, args.roi, n, np.random.default_rng(0))
    stack = analysis.FrameStack.from_device(frames, torch.from_numpy(cams).cuda(), row0=args.row0)
    res = analysis.InterpolationResult("rbf", rows, cols, analysis.light_axis(), row0=args.row0, stack=stack,
                                       precision="fast", path="rbf: banded")
    assert res.banded
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    tiles, checksum, fp64_pixels = 0, 0, 0
    for r0, tr, d_tile in analysis.table_tiles(res):
        tiles += 1
        fp64_pixels += 0 if res._band[1][3] is None else int(res._band[1][3].numel())
        checksum += int(d_tile[::997, ::4099].sum().item())          # touch the tile (and synchronise on it)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    px = rows * cols
    out = {"workload": "%d x %d pixels x %d lights, linear RBF (per-pixel lights) fit + 100x100 int32 table, banded, "
                       "one GPU, complete run" % (rows, cols, n),
           "seconds": dt, "us_per_pixel": dt * 1e6 / px, "mpixel_lights_per_s": px * n / dt / 1e6,
           "table_tiles": tiles, "table_bytes_produced": px * 10000 * 4, "pixels_routed_to_fp64": fp64_pixels,
           "checksum": checksum}
    # three pixels of the last band against scipy's algorithm in fp64
    import rti_oracle as orc
    (b0, btr), (lx, ly, w, _) = res._band
    ax = orc.light_axis()
    worst = 0.0
    for q in (0, (btr * cols) // 2 + 17, btr * cols - 1):
        l1, l2 = lx[q].cpu().numpy(), ly[q].cpu().numpy()
        dd = frames[:, b0 + q // cols, q % cols].cpu().numpy().astype(np.int32)
        ref = orc.rbf_grid(l1, l2, orc.rbf_fit(l1, l2, dd), ax)
        got = orc.rbf_grid(l1, l2, w[q].cpu().numpy(), ax)
        worst = max(worst, float(np.abs(got - ref).max() / np.abs(ref).max()))
    out["grid_max_rel_vs_oracle"] = worst
    print(json.dumps(out))
    if args.json:
        with open(os.path.join(ROOT, args.json), "w") as f:
            json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
