set -x
mkdir -p gpurun_out
N=${N:-8}
TAG=${TAG:-r2h}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29513 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline --no-parity --trace gpurun_out/${TAG}_trace_c2_n${N}_fused > gpurun_out/${TAG}_bench_c2_n${N}.json 2> gpurun_out/${TAG}_bench_c2_n${N}.err; echo "bench c2 rc=$?"
timeout 300 $TR --master-port 29514 bench.py --gpus $N --steps 20 --warmup 5 --tp-reduce nccl --no-parity --no-cpu-baseline --trace gpurun_out/${TAG}_trace_c2_n${N}_nccl > gpurun_out/${TAG}_bench_c2_n${N}_nccl.json 2> gpurun_out/${TAG}_bench_c2_n${N}_nccl.err; echo "bench c2 nccl rc=$?"
tail -n 3 gpurun_out/${TAG}_bench_c2_n${N}.err; ls -la gpurun_out | tail; true
