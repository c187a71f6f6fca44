# Second final 2-GPU check of round 2: real-rank parity tests and the bench exactly as the driver launches it
set -x
mkdir -p gpurun_out
N=${N:-2}
T=r2f2
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 900 python -m pytest tests/test_tp_gpu.py -m gpu -q > gpurun_out/${T}_tp${N}_pytest.log 2>&1; echo "pytest tp rc=$?" | tee -a gpurun_out/${T}_tp${N}_pytest.log
timeout 300 $TR --master-port 29513 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/${T}_bench_c2_n${N}.json 2> gpurun_out/${T}_bench_c2_n${N}.err; echo "bench c2 rc=$?"
timeout 300 $TR --master-port 29514 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/${T}_bench_reference_arm_n${N}.json 2> gpurun_out/${T}_bench_reference_arm_n${N}.err; echo "ref arm rc=$?"
tail -n 3 gpurun_out/${T}_tp${N}_pytest.log; tail -n 3 gpurun_out/${T}_bench_c2_n${N}.err
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2f2_bench_*_n2.json')):
    for l in open(f):
        if l.startswith('{'):
            d=json.loads(l)
            print(f.split('/')[-1], d.get('n_gpus'), round(d['ms_per_step'],2) if 'ms_per_step' in d else None, round(d.get('value',0)), (d.get('parity') or {}).get('ok'), json.dumps(d.get('e2e'))[:160])
P
