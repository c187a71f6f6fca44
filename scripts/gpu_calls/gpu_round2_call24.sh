set -x
mkdir -p gpurun_out
T=r2ah
timeout 1200 python -m pytest tests -m gpu -q --maxfail=10 > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 5 gpurun_out/${T}_pytest_gpu.log
timeout 300 python scripts/prof_host_overhead.py > gpurun_out/${T}_host_overhead.txt 2>&1; echo rc=$?; grep "^==" gpurun_out/${T}_host_overhead.txt
timeout 300 python scripts/bench_token_ops.py > gpurun_out/${T}_token_ops.jsonl 2>/dev/null; echo "tok rc=$?"
python - <<'P'
import json
rows=[json.loads(l) for l in open('gpurun_out/r2ah_token_ops.jsonl') if l.startswith('{')]
tab={}
for r in rows: tab.setdefault(r['op'],{})[r.get('config','2^25')]=r.get('frac', r.get('us'))
for k,v in tab.items(): print(f"{k[:58]:58s}", v)
P
