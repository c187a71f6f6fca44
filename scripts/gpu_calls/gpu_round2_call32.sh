# round 2, call 32: the 3072 MiB default dlogits budget on one GPU (lm-head parity tests + the headline line)
set -x
mkdir -p gpurun_out
T=r2ap
timeout 600 python -m pytest tests -m gpu -q -k "lce or config1 or guard" > gpurun_out/${T}_pytest_lce.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/${T}_pytest_lce.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-token-ops > gpurun_out/${T}_bench_c2_ws3072.json 2>/dev/null; echo "rc=$?"
python - <<'P'
import json
for l in open('gpurun_out/r2ap_bench_c2_ws3072.json'):
    if l.startswith('{'):
        d=json.loads(l); print(round(d['ms_per_step'],2), 'e2e', round(d['e2e']['ms_per_step'],2), d['parity']['ok'], d['memory'])
P
