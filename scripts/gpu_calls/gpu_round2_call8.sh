set -x
mkdir -p gpurun_out
T=r2l
timeout 600 python -m pytest tests/test_lce_gpu.py tests/test_guard_bands_gpu.py -m gpu -q -x > gpurun_out/${T}_pytest_lce.log 2>&1; rc=$?; echo "pytest lce rc=$rc"; tail -n 5 gpurun_out/${T}_pytest_lce.log
for W in 3 1 2 0 3; do
VPG_GEMM_WIDE=$W timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-parity > gpurun_out/${T}_bench_c2_wide$W.json 2> gpurun_out/${T}_bench_c2_wide$W.err; echo "bench wide=$W rc=$?"
python scripts/show_bench.py gpurun_out/${T}_bench_c2_wide$W.json | head -2
done
true
