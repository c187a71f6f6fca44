#!/bin/bash
# usage: scripts/e2e_round.sh N tag  -> bench at N GPUs with the copy-engine and the NCCL assembly of the replicated hidden
N=${1:-2}; tag=${2:-e2e}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
for g in ce nccl; do
  timeout 600 bash -c "$(declare -f run); N=$N; run 2952${#g} bench.py --gpus $N --steps 5 --warmup 3 --e2e-gather $g" \
    > gpurun_out/${tag}_${g}.json 2> gpurun_out/${tag}_${g}.err; echo "bench $g rc=$?"
  tail -3 gpurun_out/${tag}_${g}.err
done
python - <<PY
import json
for m in ("ce", "nccl"):
    try:
        d = json.loads(open("gpurun_out/${tag}_%s.json" % m).read().strip().splitlines()[-1])
        print(m, "ms/step", round(d["ms_per_step"], 3), "e2e ms", round(d["e2e"]["ms_per_step"], 3), d["e2e"]["how"][-120:])
    except Exception as e:
        print(m, "no result", e)
PY
