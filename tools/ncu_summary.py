#!/usr/bin/env python
"""Transpose `ncu -i X.ncu-rep --page raw --csv` into metric,unit,value lines (one file per
kernel launch) keeping the metrics the roofline discussion uses.

    ncu -i gpurun_out/x.ncu-rep --page raw --csv > /tmp/x.csv
    python tools/ncu_summary.py /tmp/x.csv profiles/rNN_ncu_x_raw.csv [--all]
"""
import csv
import re
import sys

KEEP = re.compile(
    r"^(Kernel Name|Block Size|Grid Size|gpu__time_duration|dram__bytes|dram__throughput|lts__t_bytes\.sum$|"
    r"lts__t_sector_hit_rate|l1tex__t_bytes\.sum$|l1tex__data_bank_conflicts|launch__|sm__throughput|"
    r"sm__warps_active|sm__inst_executed_pipe_|sm__pipe_|smsp__inst_executed\.sum$|smsp__cycles_active\.avg$|"
    r"sm__cycles_elapsed\.max$|smsp__issue_active|smsp__average_warp|smsp__warp_issue_stalled.*_per_warp_active|"
    r"sm__sass_thread_inst_executed_op|smsp__sass_inst_executed_op|l1tex__data_pipe_lsu_wavefronts_mem_shared)")


def main():
    src, dst = sys.argv[1], sys.argv[2]
    keep_all = "--all" in sys.argv
    rows = list(csv.reader(open(src)))
    hdr, units = rows[0], rows[1]
    with open(dst, "w") as f:
        for li, vals in enumerate(rows[2:]):
            f.write("# launch %d\n" % li)
            for h, u, v in zip(hdr, units, vals):
                if keep_all or KEEP.match(h):
                    f.write("%s,%s,%s\n" % (h, u, v))


if __name__ == "__main__":
    main()
