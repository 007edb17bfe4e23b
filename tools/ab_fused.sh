# A/B timings of the fused PTM kernel variants (development aid): RTI_FUSED_KERNEL = "" (warp-specialised,
# persistent), "phased" (two-phase CTAs)
for v in ws phased; do
  echo "== $v"
  RTI_FUSED_KERNEL=$v timeout 120 python tools/prof_fused.py --rows 1024 --n 6 | tail -1
  RTI_FUSED_KERNEL=$v timeout 120 python tools/prof_fused.py --rows 1024 --n 500 | tail -1
  RTI_FUSED_KERNEL=$v timeout 120 python tools/prof_fused.py --rows 512 --n 1000 --reps 3 | tail -1
done
