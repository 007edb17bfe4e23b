# A/B of library variants built as librti_b200_<tag>.so (development aid)
cd smartphone-based-rti_b200
cp librti_b200.so /tmp/librti_main.so
for v in main "$@"; do
  if [ "$v" != main ]; then cp librti_b200_$v.so librti_b200.so; fi
  echo "== $v"
  (cd .. && python tools/prof_roi.py)
done
cp /tmp/librti_main.so librti_b200.so
