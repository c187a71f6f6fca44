# round 2, call 28: refresh of the driver-op host-to-host latencies and of the reference GPU comparator at HEAD
set -x
mkdir -p gpurun_out
T=r2al
timeout 240 python scripts/bench_driver_ops_e2e.py > gpurun_out/${T}_driver_ops_e2e.jsonl 2> gpurun_out/${T}_driver_ops_e2e.err; echo "driver e2e rc=$?"; tail -n 2 gpurun_out/${T}_driver_ops_e2e.err
timeout 200 python scripts/bench_ref_gpu.py --out gpurun_out/${T}_reference_gpu_comparator.json > gpurun_out/${T}_ref_gpu.log 2>&1; echo "ref gpu rc=$?"; tail -n 5 gpurun_out/${T}_ref_gpu.log
python - <<'P'
import json
for l in open('gpurun_out/r2al_driver_ops_e2e.jsonl'):
    if l.startswith('{'):
        d=json.loads(l); print({k:(round(v,1) if isinstance(v,float) else v) for k,v in d.items()})
try: print(open('gpurun_out/r2al_reference_gpu_comparator.json').read()[:1500])
except Exception as e: print(e)
P
