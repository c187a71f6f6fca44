set -x
mkdir -p gpurun_out
T=r2ai
timeout 1200 python -m pytest tests -m gpu -q --maxfail=10 -k "rmpad or token_ops or estimators or driver_ops or actor or vendored or guard or seqlen or megatron" > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/${T}_pytest_gpu.log
timeout 300 python scripts/prof_host_overhead.py > gpurun_out/${T}_host_overhead.txt 2>&1; echo rc=$?; grep "^==" gpurun_out/${T}_host_overhead.txt
