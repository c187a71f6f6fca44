set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --maxfail=8 > gpurun_out/r2e_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2e_pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 --token-ops > gpurun_out/r2e_bench_c2_tokenops.json 2> gpurun_out/r2e_bench_c2_tokenops.err; echo "bench rc=$?"
timeout 600 python bench.py --steps 3 --warmup 3 --config c4 --no-cpu-baseline > gpurun_out/r2e_bench_c4.json 2> gpurun_out/r2e_bench_c4.err; echo "bench c4 rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2e_ncu_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-parity > gpurun_out/r2e_ncu_launches.log 2>&1; echo "ncu rc=$?"
tail -n 12 gpurun_out/r2e_pytest_gpu.log; tail -n 3 gpurun_out/r2e_bench_c2_tokenops.err; true
