set -x
mkdir -p gpurun_out
T=r2w
timeout 900 python -m pytest tests -m gpu -q --maxfail=8 -k "token_ops or estimators or driver_ops or config1 or vendored or actor or guard" > gpurun_out/${T}_pytest_tok.log 2>&1; echo "pytest rc=$?"; tail -n 5 gpurun_out/${T}_pytest_tok.log
timeout 300 python scripts/bench_token_ops.py > gpurun_out/${T}_token_ops.jsonl 2> gpurun_out/${T}_token_ops.err; echo "tok rc=$?"
grep -v '"config": "C' gpurun_out/${T}_token_ops.jsonl | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['op'][:50], d['ms'], d['frac'])"
