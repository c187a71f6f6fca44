set -x
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q --maxfail=8 > gpurun_out/r2b_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2b_pytest_gpu.log
tail -5 gpurun_out/r2b_pytest_gpu.log
timeout 300 python bench.py --steps 10 --warmup 3 --token-ops > gpurun_out/r2b_bench_c2.json 2> gpurun_out/r2b_bench_c2.err; echo "bench rc=$?"
VPG_DS_TMA=0 timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2b_bench_c2_ds_direct.json 2> gpurun_out/r2b_bench_c2_ds_direct.err
timeout 300 python bench.py --steps 10 --warmup 3 --bwd-ws-mb 24576 --no-cpu-baseline > gpurun_out/r2b_bench_c2_1chunk.json 2> gpurun_out/r2b_bench_c2_1chunk.err
timeout 300 python bench.py --config c1 --steps 10 --warmup 3 > gpurun_out/r2b_bench_c1.json 2> gpurun_out/r2b_bench_c1.err
timeout 300 python bench.py --config c5 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2b_bench_c5_n1.json 2> gpurun_out/r2b_bench_c5_n1.err
for tma in 1 0; do
 VPG_DS_TMA=$tma timeout 120 python scripts/shard_kernel_times.py --tokens 65536 --hidden 3584 --vocab-local 19008 >> gpurun_out/r2b_shard_times.jsonl 2>> gpurun_out/r2b_shard_times.err
 VPG_DS_TMA=$tma timeout 120 python scripts/shard_kernel_times.py --tokens 20480 --hidden 896 --vocab-local 151936 >> gpurun_out/r2b_shard_times.jsonl 2>> gpurun_out/r2b_shard_times.err
done
timeout 200 python scripts/shard_kernel_times.py --tokens 262144 --hidden 5120 --vocab-local 19008 --iters 3 >> gpurun_out/r2b_shard_times.jsonl 2>> gpurun_out/r2b_shard_times.err
timeout 120 python scripts/shard_kernel_times.py --tokens 65536 --hidden 8192 --vocab-local 16032 >> gpurun_out/r2b_shard_times.jsonl 2>> gpurun_out/r2b_shard_times.err
timeout 900 python scripts/run_ref_lce_test.py --mode ours > gpurun_out/r2b_ref_test_ours.log 2>&1; echo "ref test ours rc=$?" | tee -a gpurun_out/r2b_ref_test_ours.log
timeout 900 python scripts/run_ref_lce_test.py --mode reference > gpurun_out/r2b_ref_test_reference.log 2>&1; echo "ref test reference rc=$?" | tee -a gpurun_out/r2b_ref_test_reference.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2b_smoke.log 2>&1; echo "smoke rc=$?"
