set -x
mkdir -p gpurun_out
T=r2ad
for coop in 1 0; do
VPG_TOK_COOP=$coop timeout 600 python bench.py --config c3 --steps 5 --warmup 3 --no-cpu-baseline --no-token-ops --no-parity > gpurun_out/${T}_bench_c3_coop${coop}.json 2> gpurun_out/${T}_bench_c3_coop${coop}.err; echo "c3 coop$coop rc=$?"
done
VPG_TOK_COOP=1 timeout 600 python bench.py --config c3 --steps 5 --warmup 3 --no-cpu-baseline --no-token-ops > gpurun_out/${T}_bench_c3_coop1_parity.json 2>/dev/null; echo "rc=$?"
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2ad_bench_*.json')):
    for l in open(f):
        if l.startswith('{'):
            d=json.loads(l); print(f.split('/')[-1], round(d['ms_per_step'],2), 'e2e', round(d['e2e']['ms_per_step'],2), d['gpu_launches'], d['clocks']['sm_mhz'], sum(v['ms_per_step'] for v in d['roofline_kernels'].values()))
P
