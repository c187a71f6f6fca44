#!/bin/bash
# One GPU-box pass: parity tests, smoke, both bench arms, then the two ncu passes (launch list of the bench command,
# full-section capture of the four tcgen05 lm-head kernels).  Everything lands in gpurun_out/<tag>_*.
# usage: scripts/gpu_round.sh <tag>
tag=${1:-r1}
out=gpurun_out
mkdir -p $out
python -m pytest tests -m gpu -x -q > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $out/${tag}_pytest_gpu.log
python __graft_entry__.py smoke > $out/${tag}_smoke.log 2>&1; echo "smoke rc=$?"
python bench.py --steps 5 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > $out/${tag}_bench_ref.json 2> $out/${tag}_bench_ref.err; echo "ref rc=$?"
# ncu pass 1: launch list of the same bench command (cold-cache, serialised: compare shares, not absolutes)
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file $out/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu_launch.log 2>&1; echo "ncu launch rc=$?"
# ncu pass 2: full sections for the lm-head kernels of one fwd+bwd (first 4 tc_gemm launches = fwd, dlogits, dW, dH)
ncu --set full --clock-control none --import-source on -k regex:tc_gemm -c 4 -o $out/${tag}_lce_full -f \
    python scripts/prof_step.py > $out/${tag}_ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la $out | tail -12
