#!/usr/bin/env python
"""Aggregate the warp-stall samples of an `ncu --page source --csv` dump by CUDA source line.

    cuobjdump -xelf all lib.so                      # -> <file>.sm_100a.cubin
    ncu -i x.ncu-rep --page source --csv > x_src.csv
    python tools/ncu_lines.py x_src.csv rbf.sm_100a.cubin rbf_fit_kernel [top]

The SASS rows of the ncu dump are matched by order with `nvdisasm --print-line-info` of the
same kernel, which carries the `-lineinfo` file:line annotations.
"""
import csv
import re
import subprocess
import sys
from collections import defaultdict


def sass_lines(cubin, kernel):
    out = subprocess.run(["nvdisasm", "--print-line-info", cubin], capture_output=True, text=True).stdout
    cur_fn, line, res, active = None, None, [], False
    for ln in out.split("\n"):
        m = re.match(r"\s*\.text\.(\S+):", ln)
        if m:
            active = kernel in m.group(1)
            continue
        if not active:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            line = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", ln):
            res.append(line)
    return res


def main():
    src, cubin, kernel = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    rows = list(csv.reader(open(src)))
    hdr = rows[1]
    i_s = hdr.index("# Samples")
    i_x = hdr.index("Instructions Executed")
    stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_")]
    body = rows[2:]
    lines = sass_lines(cubin, kernel)
    if len(lines) != len(body):
        print("warning: %d SASS rows in ncu vs %d in nvdisasm" % (len(body), len(lines)))
    agg = defaultdict(lambda: [0, 0, defaultdict(int)])
    total = 0
    for r, l in zip(body, lines):
        s = int(r[i_s] or 0)
        a = agg[l]
        a[0] += s
        a[1] += int(r[i_x] or 0)
        for c in stall_cols:
            v = int(r[c] or 0)
            if v:
                a[2][hdr[c]] += v
        total += s
    print("total samples", total)
    for l, (s, x, st) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        tops = ", ".join("%s %d" % (k[6:], v) for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:3])
        print("%6.2f%%  %-22s inst %-10d %s" % (100.0 * s / max(total, 1), "%s:%s" % l if l else "?", x, tops))


if __name__ == "__main__":
    main()
