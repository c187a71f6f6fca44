set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,power.limit,clocks.max.sm --format=csv > gpurun_out/r2a_smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2a_pytest_gpu.log
tail -5 gpurun_out/r2a_pytest_gpu.log
timeout 300 python bench.py --steps 10 --warmup 3 --token-ops > gpurun_out/r2a_bench_c2.json 2> gpurun_out/r2a_bench_c2.err; echo "bench rc=$?"
timeout 300 python bench.py --steps 10 --warmup 3 --bwd-ws-mb 24576 --no-cpu-baseline > gpurun_out/r2a_bench_c2_1chunk.json 2> gpurun_out/r2a_bench_c2_1chunk.err
for c in c1 c3 c4; do timeout 300 python bench.py --config $c --steps 5 --warmup 3 > gpurun_out/r2a_bench_$c.json 2> gpurun_out/r2a_bench_$c.err; echo "bench $c rc=$?"; done
for sk in 1 0; do
 VPG_SPLITK=$sk timeout 120 python scripts/shard_kernel_times.py --tokens 65536 --hidden 3584 --vocab-local 19008 >> gpurun_out/r2a_shard_times.jsonl 2>> gpurun_out/r2a_shard_times.err
 VPG_SPLITK=$sk timeout 120 python scripts/shard_kernel_times.py --tokens 65536 --hidden 8192 --vocab-local 16032 >> gpurun_out/r2a_shard_times.jsonl 2>> gpurun_out/r2a_shard_times.err
 VPG_SPLITK=$sk timeout 200 python scripts/shard_kernel_times.py --tokens 262144 --hidden 5120 --vocab-local 19008 --iters 3 >> gpurun_out/r2a_shard_times.jsonl 2>> gpurun_out/r2a_shard_times.err
done
timeout 600 python scripts/bench_ref_gpu.py --out gpurun_out/r2a_reference_gpu_comparator.json > gpurun_out/r2a_ref_gpu.log 2>&1; echo "ref gpu rc=$?"
timeout 120 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2a_bench_reference_arm.json 2> gpurun_out/r2a_bench_reference_arm.err
timeout 900 python scripts/run_ref_lce_test.py --mode ours > gpurun_out/r2a_ref_test_ours.log 2>&1; echo "ref test ours rc=$?" | tee -a gpurun_out/r2a_ref_test_ours.log
timeout 900 python scripts/run_ref_lce_test.py --mode reference > gpurun_out/r2a_ref_test_reference.log 2>&1; echo "ref test reference rc=$?" | tee -a gpurun_out/r2a_ref_test_reference.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2a_smoke.log 2>&1; echo "smoke rc=$?"
