#!/bin/bash
# usage: scripts/overlap_round.sh N tag -> parity of every exchange / reduce mode, then bench A/B of the d_hidden all-reduce:
# serial NCCL vs overlapped under the d_weight GEMM (default NCCL CTAs, and capped with NCCL_MAX_CTAS)
N=${1:-2}; tag=${2:-ov}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
timeout 300 bash -c "$(declare -f run); N=$N; run 29611 scripts/tp_check.py" > gpurun_out/${tag}_check.log 2>&1; echo "check rc=$?"
grep -E "OK|MISMATCH|Error|error" gpurun_out/${tag}_check.log | head -12
i=0
for cfg in ${CFGS:-"nccl::" "overlap::" "overlap:8:" "overlap::8"}; do
  IFS=: read mode ctas reserve <<< "$cfg"; i=$((i+1))
  name=${mode}${ctas:+_ctas$ctas}${reserve:+_reserve$reserve}
  timeout 600 bash -c "$(declare -f run); N=$N; ${ctas:+export NCCL_MAX_CTAS=$ctas;} ${reserve:+export VPG_BWD_RESERVE_PAIRS=$reserve;} run 2962$i bench.py --gpus $N --steps 5 --warmup 3 --tp-reduce $mode" \
    > gpurun_out/${tag}_${name}.json 2> gpurun_out/${tag}_${name}.err; echo "bench $name rc=$?"
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/${tag}_${name}.json").read().strip().splitlines()[-1])
    print("${name}", "ms/step", round(d["ms_per_step"], 3), "e2e ms", round(d["e2e"]["ms_per_step"], 3), d["roofline_step"]["segments_ms"])
except Exception as e:
    print("${name}", "no result", e)
PY
done
