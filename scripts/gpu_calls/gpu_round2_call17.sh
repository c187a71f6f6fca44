# round 2, call 17: whiten status folded into the producing kernels (parity), sweeps of the scan block size and of the
# blocks-per-SM target of the row-wise token kernels
set -x
mkdir -p gpurun_out
T=r2z
timeout 900 python -m pytest tests -m gpu -q --maxfail=10 -k "token_ops or estimators or driver_ops or vendored or guard or config1" > gpurun_out/${T}_pytest_tok.log 2>&1; echo "pytest rc=$?"; tail -n 5 gpurun_out/${T}_pytest_tok.log
for th in 512 256 128 64; do
  VPG_SCAN_THREADS=$th timeout 300 python scripts/bench_token_ops.py --only "gae,reinforce,remax,masked_whiten" > gpurun_out/${T}_scan_${th}.jsonl 2>/dev/null; echo "scan $th rc=$?"
done
for b in 8 16 32; do
  VPG_TOK_BLOCKS_PER_SM=$b timeout 300 python scripts/bench_token_ops.py --no-configs > gpurun_out/${T}_bps_${b}.jsonl 2>/dev/null; echo "bps $b rc=$?"
done
python - <<'P'
import json
def load(p):
    try: return [json.loads(l) for l in open(p) if l.startswith('{')]
    except Exception: return []
print("scan threads: frac at 2^25 | us at C1 / C3 / C5")
rows={}
for th in (512,256,128,64):
    for r in load(f'gpurun_out/r2z_scan_{th}.jsonl'):
        key=r['op'][:40]
        rows.setdefault(key,{}).setdefault(th,[]).append(r.get('frac', r.get('us')))
for k,v in rows.items(): print(f"{k:40s}", {th: v[th] for th in v})
print("blocks per SM: frac")
tab={}
for b in (8,16,32):
    for r in load(f'gpurun_out/r2z_bps_{b}.jsonl'): tab.setdefault(r['op'][:56],{})[b]=r['frac']
for k,v in tab.items(): print(f"{k:56s}", v)
P
