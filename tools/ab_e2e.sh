# Transfer forms of rti_ptm_compute_host on one B200 (run under gpurun from the repo root):
#   the host-pipeline parity test in every form, then tools/bench_e2e.py
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "host_pipeline" > gpurun_out/e2e_test.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/e2e_test.log
timeout 400 python tools/bench_e2e.py --out gpurun_out/e2e_transfer.json > gpurun_out/e2e_transfer.log 2>&1; echo "bench_e2e rc=$?"
cat gpurun_out/e2e_transfer.log | tail -12
