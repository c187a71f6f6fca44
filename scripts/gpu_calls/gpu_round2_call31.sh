# round 2, call 31: 2-GPU bench as the driver launches it, with the 2048 MiB default dlogits budget (5 vocab chunks per rank)
set -x
mkdir -p gpurun_out
T=r2ao
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 200 $TR --master-port 29513 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/${T}_bench_c2_n2.json 2> gpurun_out/${T}_bench_c2_n2.err; echo "bench c2 rc=$?"
python - <<'P'
import json
for l in open('gpurun_out/r2ao_bench_c2_n2.json'):
    if l.startswith('{'):
        d=json.loads(l); print(d['n_gpus'], round(d['ms_per_step'],2), round(d['value']), 'e2e', round(d['e2e']['ms_per_step'],2), d['parity'], d['memory']['bwd_chunks'], d['tp_reduce'])
P
tail -n 3 gpurun_out/${T}_bench_c2_n2.err
