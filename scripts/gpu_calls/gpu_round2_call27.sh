# round 2, call 27: token ops captured into CUDA graphs (parity of replay, replay latency at the config sizes)
set -x
mkdir -p gpurun_out
T=r2ak
timeout 600 python -m pytest tests/test_token_ops_gpu.py -m gpu -q -k "graph" > gpurun_out/${T}_pytest_graph.log 2>&1; echo "pytest rc=$?"; tail -n 25 gpurun_out/${T}_pytest_graph.log
timeout 300 python scripts/bench_token_ops.py --graphs > gpurun_out/${T}_token_ops_graphs.jsonl 2>gpurun_out/${T}_token_ops_graphs.err; echo "tok rc=$?"; tail -n 3 gpurun_out/${T}_token_ops_graphs.err
python - <<'P'
import json
rows=[json.loads(l) for l in open('gpurun_out/r2ak_token_ops_graphs.jsonl') if l.startswith('{') and '"config"' in l]
tab={}
for r in rows: tab.setdefault(r['op'],{})[r['config']]=(r['us'], r.get('us_graph_replay'))
for k,v in tab.items(): print(f"{k[:58]:58s}", v)
P
