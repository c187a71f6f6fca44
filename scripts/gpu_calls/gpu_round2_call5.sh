set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --maxfail=8 > gpurun_out/r2f_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2f_pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 --token-ops > gpurun_out/r2f_bench_c2_tokenops.json 2> gpurun_out/r2f_bench_c2_tokenops.err; echo "bench rc=$?"
tail -n 12 gpurun_out/r2f_pytest_gpu.log; tail -n 3 gpurun_out/r2f_bench_c2_tokenops.err; true
