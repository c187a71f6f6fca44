# round 2, call 29: the headline configuration at smaller dlogits budgets (the reference's default split is 1.25 GB)
set -x
mkdir -p gpurun_out
T=r2am
for mb in 4096 2048 1280; do
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-token-ops --no-parity --bwd-ws-mb $mb > gpurun_out/${T}_bench_c2_ws${mb}.json 2>/dev/null; echo "rc=$?"
done
python - <<'P'
import json,glob
for mb in (4096,2048,1280):
    for l in open(f'gpurun_out/r2am_bench_c2_ws{mb}.json'):
        if l.startswith('{'):
            d=json.loads(l); print(mb, round(d['ms_per_step'],2), 'e2e', round(d['e2e']['ms_per_step'],2), d['memory']['bwd_chunks'], d['memory']['bwd_workspace_bytes'], d['memory']['peak_allocated_bytes_in_timed_region'], d['energy']['joules_per_step'])
P
