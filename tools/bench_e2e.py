#!/usr/bin/env python
"""rti_ptm_compute_host alone (bench.py's e2e leg) in its transfer forms: int32 tiles straight into the
caller's table ("wide") against int16 over PCIe + host widening threads ("narrow"), thread counts, and a
pageable destination.  Config 3: 1024 x 1024 ROI x 500 lights, 100 x 100 grid (42 GB table).

    python tools/bench_e2e.py [--rows 1024] [--out profiles/rNN_e2e_transfer.json]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import rti_b200  # noqa: E402
from rti_b200 import _capi, synthetic  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=1024)
    ap.add_argument("--cols", type=int, default=1024)
    ap.add_argument("--n", type=int, default=500)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--out", default=None)
    ap.add_argument("--modes", default="wide,narrow:16,mixed:16,mixed:12,auto,narrow:16:pageable")
    args = ap.parse_args()
    torch.cuda.set_device(0)
    lib = _capi.load()
    rows, cols, n = args.rows, args.cols, args.n
    p = rows * cols
    frames, cams = synthetic.random_stack(rows, cols, n, seed=0)
    axis = rti_b200.light_axis(0.02)
    g = axis.size
    h_frames = torch.from_numpy(frames).pin_memory()
    t0 = time.perf_counter()
    h_pinned = torch.empty((g, g, p), dtype=torch.int32, pin_memory=True)
    pin_s = time.perf_counter() - t0
    h_pageable = None
    out = []
    first = None
    for mode in args.modes.split(","):
        parts = mode.split(":")
        for name in ("RTI_HOST_TRANSFER", "RTI_HOST_THREADS", "RTI_HOST_MIX_EVERY"):
            os.environ.pop(name, None)
        for x in parts[2:]:
            if x.startswith("mix"):                               # mixed:16:mix8 = every 8th tile as int32
                os.environ["RTI_HOST_MIX_EVERY"] = x[3:]
        if parts[0] != "auto":
            os.environ["RTI_HOST_TRANSFER"] = parts[0]
        if len(parts) > 1:
            os.environ["RTI_HOST_THREADS"] = parts[1]
        pageable = "pageable" in parts[2:]
        tile_rows = ([int(x[2:]) for x in parts[2:] if x.startswith("tr")] or [0])[0]   # e.g. narrow:16:tr2
        if pageable and h_pageable is None:
            h_pageable = torch.empty((g, g, p), dtype=torch.int32)
            h_pageable.fill_(0)                                   # touch the pages
        h_table = h_pageable if pageable else h_pinned

        def step():
            _capi.call("rti_ptm_compute_host", h_frames.data_ptr(), cams.ctypes.data, n, rows, cols, 0, rows,
                       axis.ctypes.data, g, h_table.data_ptr(), None, tile_rows)
        step()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step()
        dt = (time.perf_counter() - t0) / args.steps
        moved = int(lib.rti_host_d2h_bytes())
        # a strided checksum of the table: every form must deliver the same int32 table
        chk = int(h_table.view(-1)[::9973].to(torch.int64).sum().item())
        first = chk if first is None else first
        line = {"mode": mode, "ms_per_step": dt * 1e3, "mpixel_lights_per_s": p * n / dt / 1e6,
                "d2h_bytes": moved, "pcie_gbs": moved / dt / 1e9, "table_gbs_delivered": 4.0 * g * g * p / dt / 1e9,
                "checksum_equal": chk == first}
        out.append(line)
        print(json.dumps(line), flush=True)
        lib.rti_destroy()
    print(json.dumps({"pin_42GB_s": pin_s, "cores": len(os.sched_getaffinity(0))}))
    if args.out:
        with open(os.path.join(ROOT, args.out), "w") as f:
            json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
