# A/B of library variants built as librti_b200_<tag>.so on the fit kernels (development aid):
#   bash tools/ab_fit.sh widen [more tags]
cd smartphone-based-rti_b200
cp librti_b200.so /tmp/librti_main.so
for v in main "$@"; do
  if [ "$v" != main ]; then cp librti_b200_$v.so librti_b200.so; fi
  echo "== $v"
  (cd .. && python tools/prof_ptm.py | tail -1 && python tools/prof_fused.py --rows 1024 --n 500 | tail -1 && python tools/prof_fused.py --rows 512 --n 1000 | tail -1)
  if [ "$v" != main ]; then (cd .. && python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3); fi
done
cp /tmp/librti_main.so librti_b200.so
