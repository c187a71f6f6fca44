# round 2, call 26: last full single-GPU check of HEAD (parity suite, smoke, default bench line, c3)
set -x
mkdir -p gpurun_out
T=r2aj
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${T}_pytest_gpu.log; tail -n 4 gpurun_out/${T}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/${T}_smoke.log; tail -n 2 gpurun_out/${T}_smoke.log
timeout 600 python bench.py > gpurun_out/${T}_bench_c2_default_flags.json 2> gpurun_out/${T}_bench_c2_default_flags.err; echo "bench default rc=$?"
timeout 600 python bench.py --config c3 --steps 5 --warmup 3 --no-cpu-baseline --no-token-ops > gpurun_out/${T}_bench_c3_n1.json 2> gpurun_out/${T}_bench_c3_n1.err; echo "bench c3 rc=$?"
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2aj_bench_*.json')):
    for l in open(f):
        if l.startswith('{'):
            d=json.loads(l); print(f.split('/')[-1], round(d['ms_per_step'],2), round(d['value']), 'e2e', round(d['e2e']['ms_per_step'],2), 'parity', (d.get('parity') or {}).get('ok'), 'frac', d['roofline']['frac'], 'launches', d['gpu_launches'], 'extra' in d)
P
