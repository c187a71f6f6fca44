set -x
mkdir -p gpurun_out
T=r2af
for st in 3 5 8; do
timeout 600 python bench.py --config c3 --steps $st --warmup 3 --no-cpu-baseline --no-token-ops --no-parity > gpurun_out/${T}_bench_c3_steps${st}.json 2>/dev/null; echo "rc=$?"
done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2af_bench_*.json')):
    for l in open(f):
        if l.startswith('{'):
            d=json.loads(l); print(f.split('/')[-1], round(d['ms_per_step'],2), 'e2e', round(d['e2e']['ms_per_step'],2), d['gpu_launches'], d['clocks']['sm_mhz'], round(sum(v['ms_per_step'] for v in d['roofline_kernels'].values()),1))
P
