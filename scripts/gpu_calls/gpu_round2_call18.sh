# round 2, call 18: cooperative single-pass token-mean loss (parity incl. weighted masks, A/B against the two-launch path),
# new grid / scan-block defaults
set -x
mkdir -p gpurun_out
T=r2aa
timeout 900 python -m pytest tests -m gpu -q --maxfail=10 -k "token_ops or estimators or driver_ops or vendored or guard or config1 or actor" > gpurun_out/${T}_pytest_tok.log 2>&1; echo "pytest rc=$?"; tail -n 15 gpurun_out/${T}_pytest_tok.log
VPG_TOK_COOP=0 timeout 600 python -m pytest tests -m gpu -q --maxfail=10 -k "token_ops" > gpurun_out/${T}_pytest_tok_coop0.log 2>&1; echo "pytest coop0 rc=$?"; tail -n 3 gpurun_out/${T}_pytest_tok_coop0.log
timeout 300 python scripts/bench_token_ops.py > gpurun_out/${T}_token_ops.jsonl 2>gpurun_out/${T}_token_ops.err; echo "tok rc=$?"; tail -n 3 gpurun_out/${T}_token_ops.err
VPG_TOK_COOP=0 timeout 300 python scripts/bench_token_ops.py --only ppo_loss > gpurun_out/${T}_token_ops_coop0.jsonl 2>/dev/null; echo "tok coop0 rc=$?"
python - <<'P'
import json
def load(p):
    try: return [json.loads(l) for l in open(p) if l.startswith('{')]
    except Exception: return []
for r in load('gpurun_out/r2aa_token_ops.jsonl'):
    print(f"{r.get('config','2^25'):12s} {r['op'][:58]:58s}", r.get('ms', r.get('us')), r.get('frac',''))
print('--- coop0')
for r in load('gpurun_out/r2aa_token_ops_coop0.jsonl'):
    print(f"{r.get('config','2^25'):12s} {r['op'][:58]:58s}", r.get('ms', r.get('us')), r.get('frac',''))
P
