#!/usr/bin/env bash
# One parameterised GPU validation script (replaces the per-call scratch files of rounds 1-2, which stay in git history).
# Run it on the GPU box through gpurun, e.g.
#   gpurun --timeout 1500 -- 'TAG=r3 bash scripts/gpu_check.sh suite smoke bench'
#   gpurun --gpus 2 --timeout 1500 -- 'TAG=r3 N=2 bash scripts/gpu_check.sh tp bench'
#   gpurun --timeout 1500 -- 'TAG=r3 bash scripts/gpu_check.sh ncu'
# Stages (any subset, in the order given):
#   suite   pytest -m gpu (the whole GPU parity suite)
#   smoke   __graft_entry__.smoke()
#   ref     bench.py --impl reference (the reference's CPU arm)
#   bench   bench.py at N GPUs with default flags (headline + extra.configs + verdict)
#   configs every BASELINE config on its own, 5 steps each (N=1: c1 c3 c4 c5; N>1: c4 c5)
#   tokops  token-op HBM roofline table
#   tp      N>1: tp_check.py (all exchange / reduction modes vs fp32 torch) and the missing-peer timeout check
#   ncu     ncu launch list of the bench command + one --set full capture of the lm-head kernels + token-kernel table
# Everything lands under gpurun_out/${TAG}_*; copy what should be judged into profiles/.
set -u
T=${TAG:-r3}
N=${N:-1}
STEPS=${STEPS:-20}
out=gpurun_out
mkdir -p $out
run_bench() {  # args: tag-suffix, bench flags...
    local name=$1; shift
    if [ "$N" = "1" ]; then
        timeout 900 python bench.py "$@" > $out/${T}_${name}_n1.json 2> $out/${T}_${name}_n1.err
    else
        timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
            --master-port ${PORT:-29620} bench.py --gpus $N "$@" > $out/${T}_${name}_n${N}.json 2> $out/${T}_${name}_n${N}.err
    fi
    echo "bench $name N=$N rc=$?"
}
for stage in "$@"; do
    case $stage in
    suite)
        timeout 1500 python -m pytest tests -m gpu -q -rs > $out/${T}_pytest_gpu_n${N}.log 2>&1
        echo "pytest rc=$?" | tee -a $out/${T}_pytest_gpu_n${N}.log; tail -n 4 $out/${T}_pytest_gpu_n${N}.log ;;
    smoke)
        timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $out/${T}_smoke.log 2>&1
        echo "smoke rc=$?" | tee -a $out/${T}_smoke.log; tail -n 2 $out/${T}_smoke.log ;;
    ref)
        timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $out/${T}_bench_reference_arm.json \
            2> $out/${T}_bench_reference_arm.err; echo "reference arm rc=$?" ;;
    bench)
        run_bench bench_default --steps $STEPS --warmup 5 ;;
    configs)
        cfgs="c1 c3 c4 c5"; [ "$N" != "1" ] && cfgs="c4 c5"
        for C in $cfgs; do
            run_bench bench_$C --config $C --steps 5 --warmup 3 --no-cpu-baseline --no-token-ops
        done ;;
    tokops)
        timeout 300 python scripts/bench_token_ops.py > $out/${T}_token_ops_hbm_roofline.jsonl 2> $out/${T}_token_ops.err
        echo "token ops rc=$?" ;;
    tp)
        timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
            --master-port 29511 scripts/tp_check.py > $out/${T}_tp${N}_check.log 2>&1; echo "tp_check rc=$?"
        VPG_AR_TIMEOUT_MS=1500 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 \
            --master-addr 127.0.0.1 --master-port 29513 scripts/tp_timeout_check.py > $out/${T}_tp2_timeout.log 2>&1
        echo "tp_timeout_check rc=$?" ;;
    ncu)
        M=gpu__time_duration.sum,dram__bytes.sum.per_second,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,sm__throughput.avg.pct_of_peak_sustained_elapsed,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__grid_size,launch__block_size,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum
        timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
            --log-file $out/${T}_ncu_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-parity \
            --no-token-ops --no-extra-configs --no-reference-gpu > $out/${T}_ncu_launches.log 2>&1; echo "ncu launch list rc=$?"
        timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_gemm -c 16 -o $out/${T}_lce_full -f \
            python scripts/prof_step.py > $out/${T}_ncu_full.log 2>&1; echo "ncu full rc=$?"
        python scripts/ncu_summary.py $out/${T}_lce_full.ncu-rep $out/${T}_ncu_full_tc_gemm_kernels.csv \
            > $out/${T}_ncu_full_summary.txt 2>&1
        rm -f $out/${T}_lce_full.ncu-rep
        timeout 600 ncu --metrics $M --clock-control none \
            -k regex:"ppo_loss|mask_count|gae_kernel|whiten|broadcast_mask|row_score|data_metrics|seq_lengths|scatter_narrow|unpad|agg_loss|value_loss|reverse_scan|kl_reward|grpo_group|kl_penalty|masked_sum|permute_rows|row_broadcast" \
            --csv --log-file $out/${T}_ncu_token_ops_long.csv python scripts/prof_token_ops.py > $out/${T}_ncu_tok.log 2>&1
        echo "ncu token kernels rc=$?" ;;
    *) echo "unknown stage: $stage" ;;
    esac
done
python scripts/show_bench.py $out/${T}_bench_*.json 2>/dev/null || true
true
