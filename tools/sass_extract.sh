#!/bin/bash
# Per-kernel SASS opcode histogram of the built library: the evidence that the tcgen05 / TMEM / TMA /
# fp64-tensor-core claims of DESIGN.md are in the shipped binary (the .so itself is git-ignored).
#   tools/sass_extract.sh [round-tag]   ->  profiles/<tag>_sass_summary.txt
set -euo pipefail
cd "$(dirname "$0")/.."
TAG=${1:-r02}
SO=smartphone-based-rti_b200/librti_b200.so
OUT=profiles/${TAG}_sass_summary.txt
[ -f "$SO" ] || python __graft_entry__.py build
{
  echo "# cuobjdump -sass $SO  (nvcc $(nvcc --version | grep -o 'release [0-9.]*'), sm_100a), $(date -u +%Y-%m-%dT%H:%MZ)"
  echo "# columns: kernel, then count of each opcode family found in its SASS"
  echo "# UTCHMMA = tcgen05.mma (fp16/bf16 kinds), UTCBAR = tcgen05.commit, LDTM = tcgen05.ld (TMEM -> registers),"
  echo "# UTMALDG/UTMASTG = cp.async.bulk.tensor (TMA tile load/store), UBLKCP = cp.async.bulk (TMA 1-D),"
  echo "# SYNCS = mbarrier, DMMA = fp64 tensor core mma.sync, LDGSTS = cp.async, MUFU = special function unit;"
  echo "# 2CTA = instructions carrying the .2CTA modifier (cta_group::2: pair MMA, pair TMA load, multicast commit)"
  cuobjdump -sass "$SO" | awk '
    /Function :/ { fn=$3; order[++n]=fn; next }
    fn != "" {
      if (match($0, /(UTCHMMA|UTCQMMA|UTCIMMA|UTCBAR|UTCATOMSWS|LDTM|STTM|UTMALDG|UTMASTG|UTMAPF|UBLKCP|UBLKPF|SYNCS|DMMA|HMMA|LDGSTS|MUFU|DFMA|FFMA|STG|LDG|ATOMS|REDUX)[A-Z0-9_.]*/)) {
        full=substr($0, RSTART, RLENGTH); op=full; sub(/\..*/, "", op); c[fn SUBSEP op]++; ops[op]=1
        if (full ~ /2CTA/) { c[fn SUBSEP "2CTA"]++; ops["2CTA"]=1 }
      }
    }
    END {
      for (i=1;i<=n;i++) { fn=order[i]; line=fn;
        for (op in ops) if (c[fn SUBSEP op]) line=line sprintf("  %s=%d", op, c[fn SUBSEP op]);
        print line }
    }' | while read -r mangled rest; do printf "%s  %s\n" "$(echo "$mangled" | c++filt | cut -c1-110)" "$rest"; done | sort
} > "$OUT"
echo "wrote $OUT ($(wc -l < "$OUT") lines)"
