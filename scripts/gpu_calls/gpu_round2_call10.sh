set -x
mkdir -p gpurun_out
T=r2o
timeout 600 python scripts/sweep_fwd_splits.py > gpurun_out/${T}_sweep_fwd_splits.jsonl 2> gpurun_out/${T}_sweep_fwd_splits.err; echo "sweep rc=$?"
cat gpurun_out/${T}_sweep_fwd_splits.jsonl
timeout 300 python -m pytest tests/test_token_ops_gpu.py -m gpu -q -k grpo > gpurun_out/${T}_pytest_grpo.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/${T}_pytest_grpo.log
