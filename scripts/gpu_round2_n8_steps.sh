set -x
mkdir -p gpurun_out
N=${N:-8}
TAG=${TAG:-r2i}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29513 bench.py --gpus $N --steps 30 --warmup 5 --no-cpu-baseline --no-parity > gpurun_out/${TAG}_bench_c2_n${N}.json 2> gpurun_out/${TAG}_bench_c2_n${N}.err; echo "bench c2 rc=$?"
tail -n 3 gpurun_out/${TAG}_bench_c2_n${N}.err; true
