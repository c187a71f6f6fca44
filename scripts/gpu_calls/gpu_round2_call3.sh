set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --maxfail=8 > gpurun_out/r2d_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2d_pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 --token-ops > gpurun_out/r2d_bench_c2_tokenops.json 2> gpurun_out/r2d_bench_c2_tokenops.err; echo "bench rc=$?"
timeout 600 python scripts/bench_driver_ops_e2e.py > gpurun_out/r2d_driver_ops_e2e.jsonl 2> gpurun_out/r2d_driver_ops_e2e.err; echo "driver e2e rc=$?"
SAN_TIMEOUT=700 bash scripts/sanitize.sh memcheck
tail -n 5 gpurun_out/r2d_pytest_gpu.log; tail -n 3 gpurun_out/r2d_driver_ops_e2e.err; cat gpurun_out/r2_sanitizer_summary.txt; true
