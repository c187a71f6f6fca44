set -x
mkdir -p gpurun_out
T=r2s
timeout 600 ncu --section SpeedOfLight --section MemoryWorkloadAnalysis --section WarpStateStats --section Occupancy --section LaunchStats --clock-control none -k regex:"ppo_loss|mask_count|gae_kernel|whiten|broadcast_mask|row_score|data_metrics|scatter_narrow|unpad|agg_loss|value_loss|reverse_scan|kl_reward|grpo_group" -o /tmp/${T}_tok -f python scripts/prof_token_ops.py > gpurun_out/${T}_ncu_tok.log 2>&1; echo "ncu rc=$?"
ncu -i /tmp/${T}_tok.ncu-rep --page raw --csv > gpurun_out/${T}_tok_raw.csv 2>/dev/null
ls -la gpurun_out/${T}_* /tmp/${T}_tok.ncu-rep
