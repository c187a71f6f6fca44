# round 2, call 16: validate the loss-sweep diet (parity suite, token-op table A/B), the c3 parity probe, default bench
set -x
mkdir -p gpurun_out
T=r2y
timeout 1200 python -m pytest tests -m gpu -q --maxfail=10 > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 8 gpurun_out/${T}_pytest_gpu.log
timeout 300 python scripts/bench_token_ops.py > gpurun_out/${T}_token_ops.jsonl 2> gpurun_out/${T}_token_ops.err; echo "tok rc=$?"
VPG_TOK_SPEC=0 timeout 300 python scripts/bench_token_ops.py > gpurun_out/${T}_token_ops_spec0.jsonl 2>/dev/null; echo "tok spec0 rc=$?"
VPG_TOK_BLOCKS_PER_SM=16 timeout 300 python scripts/bench_token_ops.py > gpurun_out/${T}_token_ops_bps16.jsonl 2>/dev/null; echo "tok bps16 rc=$?"
python - <<'P'
import json
def load(p):
    try: return {json.loads(l)['op']: json.loads(l) for l in open(p) if l.startswith('{') and '"config"' not in l}
    except Exception as e: return {}
a,b,c=[load(f'gpurun_out/r2y_token_ops{s}.jsonl') for s in ('','_spec0','_bps16')]
for k in a: print(f"{k[:56]:56s} {a[k]['ms']:.4f} {a[k]['frac']:.3f} | spec0 {b.get(k,{}).get('frac')} | bps16 {c.get(k,{}).get('frac')}")
P
timeout 600 python bench.py --config c3 --steps 3 --warmup 3 --no-cpu-baseline --no-token-ops > gpurun_out/${T}_bench_c3.json 2> gpurun_out/${T}_bench_c3.err; echo "c3 rc=$?"; tail -n 3 gpurun_out/${T}_bench_c3.err
python - <<'P'
import json
for l in open('gpurun_out/r2y_bench_c3.json'):
    if l.startswith('{'):
        d=json.loads(l); print('c3', d['ms_per_step'], d['value'], json.dumps(d['parity'])[:900]); print(json.dumps(d['roofline'])[:600]); print(json.dumps(d['roofline_kernels'])[:900])
P
timeout 600 python bench.py > gpurun_out/${T}_bench_c2_default.json 2> gpurun_out/${T}_bench_c2_default.err; echo "c2 rc=$?"; tail -n 3 gpurun_out/${T}_bench_c2_default.err
python - <<'P'
import json
for l in open('gpurun_out/r2y_bench_c2_default.json'):
    if l.startswith('{'):
        d=json.loads(l); print('c2', d['ms_per_step'], d['value'], d['e2e']['ms_per_step'], d['parity']['ok'], d['clocks']); print(json.dumps(d.get('extra'))[:1500])
P
timeout 300 python __graft_entry__.py smoke > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/${T}_smoke.log
