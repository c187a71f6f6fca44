# Final single-GPU validation of the round: parity suite, smoke, both bench arms, every BASELINE config at N=1,
# the ncu launch list of the bench command and one full-section capture of the lm-head kernels.
set -x
T=${TAG:-r2q}
out=gpurun_out
mkdir -p $out
timeout 1500 python -m pytest tests -m gpu -q > $out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $out/${T}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $out/${T}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $out/${T}_smoke.log
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $out/${T}_bench_reference_arm.json 2> $out/${T}_bench_reference_arm.err; echo "ref rc=$?"
timeout 600 python bench.py > $out/${T}_bench_c2_default_flags.json 2> $out/${T}_bench_c2_default_flags.err; echo "bench default rc=$?"
timeout 600 python bench.py --steps 20 --warmup 5 --token-ops --no-cpu-baseline > $out/${T}_bench_c2_n1.json 2> $out/${T}_bench_c2_n1.err; echo "bench c2 rc=$?"
for C in c1 c3 c4 c5; do
timeout 600 python bench.py --config $C --steps 5 --warmup 3 --no-cpu-baseline > $out/${T}_bench_${C}_n1.json 2> $out/${T}_bench_${C}_n1.err; echo "bench $C rc=$?"
done
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/${T}_ncu_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-parity > $out/${T}_ncu_launches.log 2>&1; echo "ncu launch rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_gemm -c 16 -o $out/${T}_lce_full -f python scripts/prof_step.py > $out/${T}_ncu_full.log 2>&1; echo "ncu full rc=$?"
python scripts/ncu_summary.py $out/${T}_lce_full.ncu-rep $out/${T}_ncu_full_tc_gemm_kernels.csv > $out/${T}_ncu_full_summary.txt 2>&1
rm -f $out/${T}_lce_full.ncu-rep
tail -n 4 $out/${T}_pytest_gpu.log; tail -n 2 $out/${T}_smoke.log; true
