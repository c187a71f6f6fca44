# Second final single-GPU validation of round 2 (after the token-kernel work): parity suite, smoke, both bench arms, every
# BASELINE config at N=1, the token-op table + its ncu metrics, the ncu launch list of the bench command and one
# full-section capture of the lm-head kernels.
set -x
T=${TAG:-r2f2}
out=gpurun_out
mkdir -p $out
timeout 1500 python -m pytest tests -m gpu -q > $out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $out/${T}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $out/${T}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $out/${T}_smoke.log
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $out/${T}_bench_reference_arm.json 2> $out/${T}_bench_reference_arm.err; echo "ref rc=$?"
timeout 600 python bench.py > $out/${T}_bench_c2_default_flags.json 2> $out/${T}_bench_c2_default_flags.err; echo "bench default rc=$?"
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $out/${T}_bench_c2_n1.json 2> $out/${T}_bench_c2_n1.err; echo "bench c2 rc=$?"
for C in c1 c3 c4 c5; do
timeout 600 python bench.py --config $C --steps 5 --warmup 3 --no-cpu-baseline --no-token-ops > $out/${T}_bench_${C}_n1.json 2> $out/${T}_bench_${C}_n1.err; echo "bench $C rc=$?"
done
timeout 300 python scripts/bench_token_ops.py > $out/${T}_token_ops_hbm_roofline.jsonl 2> $out/${T}_token_ops.err; echo "tok rc=$?"
M=gpu__time_duration.sum,dram__bytes.sum.per_second,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,sm__throughput.avg.pct_of_peak_sustained_elapsed,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__grid_size,launch__block_size,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum
timeout 600 ncu --metrics $M --clock-control none -k regex:"ppo_loss|mask_count|gae_kernel|whiten|broadcast_mask|row_score|data_metrics|seq_lengths|scatter_narrow|unpad|agg_loss|value_loss|reverse_scan|kl_reward|grpo_group|kl_penalty|masked_sum|permute_rows|row_broadcast" --csv --log-file $out/${T}_ncu_token_ops_long.csv python scripts/prof_token_ops.py > $out/${T}_ncu_tok.log 2>&1; echo "ncu tok rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/${T}_ncu_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-parity --no-token-ops > $out/${T}_ncu_launches.log 2>&1; echo "ncu launch rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_gemm -c 16 -o $out/${T}_lce_full -f python scripts/prof_step.py > $out/${T}_ncu_full.log 2>&1; echo "ncu full rc=$?"
python scripts/ncu_summary.py $out/${T}_lce_full.ncu-rep $out/${T}_ncu_full_tc_gemm_kernels.csv > $out/${T}_ncu_full_summary.txt 2>&1
rm -f $out/${T}_lce_full.ncu-rep
tail -n 4 $out/${T}_pytest_gpu.log; tail -n 2 $out/${T}_smoke.log
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2f2_bench_*.json')):
    for l in open(f):
        if l.startswith('{'):
            d=json.loads(l)
            if 'ms_per_step' in d:
                print(f.split('/')[-1], round(d['ms_per_step'],2), round(d['value']), 'e2e', round(d['e2e']['ms_per_step'],2) if d.get('e2e') and d['e2e'].get('ms_per_step') else None, 'parity', (d.get('parity') or {}).get('ok'), 'frac', d.get('roofline',{}).get('frac'))
            else: print(f.split('/')[-1], json.dumps(d)[:300])
P
true
