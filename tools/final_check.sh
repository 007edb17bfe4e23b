# Round-end verification on one B200 (run under gpurun from the repo root): the GPU tests, smoke(), both
# bench arms, the per-kernel microbenchmarks and two ncu captures, each under its own timeout, logs in gpurun_out/.
#   gpurun --timeout 900 -- 'bash tools/final_check.sh'        (NCU=1 adds the two ncu captures, KERNELS=0 skips the microbenchmarks)
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > $O/gpu.txt 2>&1
timeout 420 python -m pytest tests -m gpu -x -q > $O/gputest.log 2>&1; echo "pytest rc=$?" | tee -a $O/gputest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/smoke.log
timeout 120 python bench.py --impl reference > $O/bench_reference_n1.json 2> $O/bench_reference_n1.err; echo "bench ref rc=$?"
timeout 400 python bench.py > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?"
[ "$KERNELS" = 0 ] || timeout 240 python tools/bench_kernels.py --out gpurun_out/kernels.json > $O/kernels.log 2>&1; echo "kernels rc=$?"
if [ "$NCU" = 1 ]; then
timeout 150 ncu --set full --clock-control none --import-source on -k regex:ptm_fit_soa_kernel -c 1 -s 1 -o $O/ncu_ptm_fit_soa -f \
    python tools/prof_ptm.py > $O/ncu_ptm_fit_soa.log 2>&1; echo "ncu fit_soa rc=$?"
timeout 150 ncu --set full --clock-control none --import-source on -k regex:relight -c 4 -o $O/ncu_relight -f \
    python tools/prof_relight.py > $O/ncu_relight.log 2>&1; echo "ncu relight rc=$?"
fi
tail -3 $O/gputest.log; tail -2 $O/smoke.log
