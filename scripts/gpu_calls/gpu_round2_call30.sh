# round 2, call 30: full single-GPU check with the 2048 MiB default dlogits budget
set -x
mkdir -p gpurun_out
T=r2an
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${T}_pytest_gpu.log; tail -n 4 gpurun_out/${T}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/${T}_smoke.log
timeout 600 python bench.py > gpurun_out/${T}_bench_c2_default_flags.json 2> gpurun_out/${T}_bench_c2_default_flags.err; echo "bench default rc=$?"
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-token-ops > gpurun_out/${T}_bench_c2_n1.json 2>/dev/null; echo "bench 20 rc=$?"
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2an_bench_*.json')):
    for l in open(f):
        if l.startswith('{'):
            d=json.loads(l); print(f.split('/')[-1], round(d['ms_per_step'],2), round(d['value']), 'e2e', round(d['e2e']['ms_per_step'],2), 'parity', (d.get('parity') or {}).get('ok'), 'frac', round(d['roofline']['frac'],3), d['memory']['bwd_chunks'], d['memory']['bwd_workspace_bytes'], d['memory']['budget'][:40])
P
