# rti_table_tile_to_host on one B200 (run under gpurun from the repo root): its own test, the tests of the
# drop-in calls that go through it, then prepare_images_data at 1024^2 both ways
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_rbf_shared.py -m gpu -x -q -k "table_tile_to_host or host_pipeline or call_surface or compute or rbf_banded or reference_golden" > gpurun_out/tth_test.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/tth_test.log
timeout 300 python tools/bench_prepare.py > gpurun_out/prepare.log 2>&1; echo "bench_prepare rc=$?"
tail -6 gpurun_out/prepare.log
