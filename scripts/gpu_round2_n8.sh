set -x
mkdir -p gpurun_out
N=${N:-8}
TAG=${TAG:-r2r}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29511 scripts/tp_check.py --time > gpurun_out/${TAG}_tp${N}_check.log 2>&1; echo "tp_check rc=$?" | tee -a gpurun_out/${TAG}_tp${N}_check.log
timeout 600 python -m pytest tests/test_tp_gpu.py -m gpu -q > gpurun_out/${TAG}_tp${N}_pytest.log 2>&1; echo "pytest tp rc=$?" | tee -a gpurun_out/${TAG}_tp${N}_pytest.log
timeout 300 $TR --master-port 29513 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${TAG}_bench_c2_n${N}.json 2> gpurun_out/${TAG}_bench_c2_n${N}.err; echo "bench c2 rc=$?"
timeout 300 $TR --master-port 29514 bench.py --gpus $N --steps 20 --warmup 5 --tp-reduce nccl --no-parity --no-cpu-baseline > gpurun_out/${TAG}_bench_c2_n${N}_nccl.json 2> gpurun_out/${TAG}_bench_c2_n${N}_nccl.err; echo "bench c2 nccl rc=$?"
timeout 300 $TR --master-port 29515 bench.py --gpus $N --config c4 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_c4_n${N}.json 2> gpurun_out/${TAG}_bench_c4_n${N}.err; echo "bench c4 rc=$?"
if [ "$N" = "8" ]; then
timeout 400 $TR --master-port 29516 bench.py --gpus $N --config c5 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_c5_n${N}.json 2> gpurun_out/${TAG}_bench_c5_n${N}.err; echo "bench c5 rc=$?"
fi
tail -n 3 gpurun_out/${TAG}_tp${N}_check.log; tail -n 3 gpurun_out/${TAG}_bench_c2_n${N}.err; tail -n 3 gpurun_out/${TAG}_tp${N}_pytest.log; true
