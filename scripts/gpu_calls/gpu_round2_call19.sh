# round 2, call 19: cooperative single-pass loss with grid-stride units
set -x
mkdir -p gpurun_out
T=r2ab
timeout 900 python -m pytest tests -m gpu -q --maxfail=10 -k "token_ops or config1 or actor or guard" > gpurun_out/${T}_pytest_tok.log 2>&1; echo "pytest rc=$?"; tail -n 5 gpurun_out/${T}_pytest_tok.log
timeout 300 python scripts/bench_token_ops.py --only ppo_loss > gpurun_out/${T}_token_ops.jsonl 2>gpurun_out/${T}_token_ops.err; echo "tok rc=$?"; tail -n 3 gpurun_out/${T}_token_ops.err
cat gpurun_out/${T}_token_ops.jsonl | cut -c1-200
