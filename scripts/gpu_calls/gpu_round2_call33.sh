# round 2, call 33: the default bench line and smoke at HEAD (3072 MiB default dlogits budget)
set -x
mkdir -p gpurun_out
T=r2aq
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/${T}_smoke.log
timeout 400 python bench.py > gpurun_out/${T}_bench_c2_default_flags.json 2> gpurun_out/${T}_bench_c2_default_flags.err; echo "bench default rc=$?"
python - <<'P'
import json
for l in open('gpurun_out/r2aq_bench_c2_default_flags.json'):
    if l.startswith('{'):
        d=json.loads(l); print(round(d['ms_per_step'],2), round(d['value']), 'e2e', round(d['e2e']['ms_per_step'],2), d['parity']['ok'], round(d['roofline']['frac'],3), d['memory']['bwd_chunks'], d['energy']['joules_per_step'], d['cpu_baseline']['value'], 'extra' in d)
P
