set -x
mkdir -p gpurun_out
timeout 600 python scripts/ab_pad_input.py > gpurun_out/r2j_ab_pad_input.jsonl 2> gpurun_out/r2j_ab_pad_input.err; echo "ab rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_gemm -c 4 -o gpurun_out/r2j_lce_full -f python scripts/prof_step.py > gpurun_out/r2j_ncu_full.log 2>&1; echo "ncu full rc=$?"
python scripts/ncu_summary.py gpurun_out/r2j_lce_full.ncu-rep gpurun_out/r2j_ncu_full_tc_gemm_kernels.csv > gpurun_out/r2j_ncu_full_summary.txt 2>&1
cat gpurun_out/r2j_ab_pad_input.jsonl; ls -la gpurun_out | tail -5; true
