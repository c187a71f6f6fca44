set -x
mkdir -p gpurun_out
N=${N:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29511 scripts/tp_check.py --time > gpurun_out/r2c_tp${N}_check.log 2>&1; echo "tp_check rc=$?" | tee -a gpurun_out/r2c_tp${N}_check.log
timeout 300 $TR --master-port 29512 scripts/tp_check.py --tokens 8192 --hidden 3584 --vocab 38016 > gpurun_out/r2c_tp${N}_check_b.log 2>&1; echo "tp_check b rc=$?" | tee -a gpurun_out/r2c_tp${N}_check_b.log
timeout 600 python -m pytest tests/test_tp_gpu.py -m gpu -q > gpurun_out/r2c_tp${N}_pytest.log 2>&1; echo "pytest tp rc=$?" | tee -a gpurun_out/r2c_tp${N}_pytest.log
timeout 300 $TR --master-port 29513 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r2c_bench_c2_n${N}.json 2> gpurun_out/r2c_bench_c2_n${N}.err; echo "bench c2 rc=$?"
timeout 300 $TR --master-port 29514 bench.py --gpus $N --steps 10 --warmup 3 --tp-reduce nccl --no-parity > gpurun_out/r2c_bench_c2_n${N}_nccl.json 2> gpurun_out/r2c_bench_c2_n${N}_nccl.err; echo "bench c2 nccl rc=$?"
timeout 300 $TR --master-port 29515 bench.py --gpus $N --config c4 --steps 5 --warmup 3 > gpurun_out/r2c_bench_c4_n${N}.json 2> gpurun_out/r2c_bench_c4_n${N}.err; echo "bench c4 rc=$?"
tail -n 3 gpurun_out/r2c_tp${N}_check.log; tail -n 3 gpurun_out/r2c_bench_c2_n${N}.err; true
