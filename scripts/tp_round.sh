#!/bin/bash
# usage: scripts/tp_round.sh N tag   -> vocab-parallel parity + bench (NVLink-fused vs NCCL exchange) on N GPUs
N=${1:-2}; tag=${2:-tp}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
timeout 300 bash -c "$(declare -f run); N=$N; run 29511 scripts/tp_check.py" > gpurun_out/${tag}_check.log 2>&1; echo "check rc=$?"
grep -E "OK|MISMATCH|multicast" gpurun_out/${tag}_check.log
timeout 600 bash -c "$(declare -f run); N=$N; run 29512 bench.py --gpus $N --steps 5 --warmup 3 --tp-exchange nvlink" > gpurun_out/${tag}_bench_nvlink.json 2> gpurun_out/${tag}_bench_nvlink.err; echo "bench nvlink rc=$?"
timeout 600 bash -c "$(declare -f run); N=$N; run 29513 bench.py --gpus $N --steps 5 --warmup 3 --tp-exchange nccl" > gpurun_out/${tag}_bench_nccl.json 2> gpurun_out/${tag}_bench_nccl.err; echo "bench nccl rc=$?"
python - <<PY
import json
for m in ("nvlink", "nccl"):
    try:
        d = json.loads(open("gpurun_out/${tag}_bench_%s.json" % m).read().strip().splitlines()[-1])
        print(m, "ms/step", round(d["ms_per_step"], 3), "e2e ms", round(d["e2e"]["ms_per_step"], 3), "tok/s", round(d["value"]),
              d["roofline_step"]["segments_ms"], d["config"]["parallelism"][:60])
    except Exception as e:
        print(m, "no result", e)
PY
