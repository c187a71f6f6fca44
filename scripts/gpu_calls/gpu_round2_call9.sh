set -x
mkdir -p gpurun_out
T=r2m
timeout 900 python -m pytest tests -m gpu -q --maxfail=8 -k "token_ops or estimators or driver_ops or config1 or rmpad or guard or vendored or actor" > gpurun_out/${T}_pytest_tok.log 2>&1; echo "pytest rc=$?"; tail -n 8 gpurun_out/${T}_pytest_tok.log
timeout 300 python scripts/bench_token_ops.py > gpurun_out/${T}_token_ops.jsonl 2> gpurun_out/${T}_token_ops.err; echo "tok rc=$?"
cat gpurun_out/${T}_token_ops.jsonl | cut -c1-200
