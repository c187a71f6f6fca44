set -x
mkdir -p gpurun_out
T=r2k
timeout 300 python -m pytest tests/test_lce_gpu.py -m gpu -q -x -k "gemm" > gpurun_out/${T}_pytest_gemm.log 2>&1; rc=$?; echo "pytest gemm rc=$rc"; tail -n 15 gpurun_out/${T}_pytest_gemm.log
if [ $rc -ne 0 ]; then exit 0; fi
timeout 1500 python -m pytest tests -m gpu -q --maxfail=8 > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${T}_pytest_gpu.log
tail -n 12 gpurun_out/${T}_pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_bench_c2_wide.json 2> gpurun_out/${T}_bench_c2_wide.err; echo "bench wide rc=$?"
VPG_GEMM_WIDE=0 timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-parity > gpurun_out/${T}_bench_c2_narrow.json 2> gpurun_out/${T}_bench_c2_narrow.err; echo "bench narrow rc=$?"
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-parity > gpurun_out/${T}_bench_c2_wide2.json 2> gpurun_out/${T}_bench_c2_wide2.err; echo "bench wide2 rc=$?"
timeout 600 python scripts/ab_pad_input.py > gpurun_out/${T}_ab_pad_input.jsonl 2> gpurun_out/${T}_ab_pad_input.err; echo "ab rc=$?"
tail -n 3 gpurun_out/${T}_bench_c2_wide.err; true
