set -x
mkdir -p gpurun_out
T=r2p
SWEEP_SPLITS=4,5,6,7,8,9,11,13 timeout 900 python scripts/sweep_fwd_splits.py c2s8 c4s8 c5s c2s4 c2s2 c4s4 c4s2 c1 c4 > gpurun_out/${T}_sweep_fwd_splits.jsonl 2> gpurun_out/${T}_sweep_fwd_splits.err; echo "sweep rc=$?"
cat gpurun_out/${T}_sweep_fwd_splits.jsonl | cut -c1-140
