set -x
mkdir -p gpurun_out
T=r2ac
M=gpu__time_duration.sum,dram__bytes.sum.per_second,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,sm__throughput.avg.pct_of_peak_sustained_elapsed,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__grid_size,launch__block_size,launch__occupancy_limit_shared_mem,launch__occupancy_limit_registers,launch__shared_mem_per_block_dynamic,launch__shared_mem_config_size,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio,smsp__average_warps_issue_stalled_membar_per_issue_active.ratio,smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct
for coop in 1 0; do
VPG_TOK_COOP=$coop ncu --metrics $M --clock-control none -k regex:"ppo_loss|mask_count|value_loss" -c 12 --csv --log-file gpurun_out/${T}_ncu_coop${coop}.csv python scripts/prof_ppo_loss.py > gpurun_out/${T}_ncu_coop${coop}.log 2>&1; echo "ncu coop$coop rc=$?"
done
python - <<'P'
import csv
for coop in (1,0):
    rows=[r for r in csv.reader(open(f'gpurun_out/r2ac_ncu_coop{coop}.csv')) if len(r)>10]
    h=rows[0]; n=h.index('Metric Name'); v=h.index('Metric Value'); i=h.index('ID'); kn=h.index('Kernel Name')
    d={}; names={}
    for r in rows[1:]:
        d.setdefault(r[i],{})[r[n]]=r[v]; names[r[i]]=r[kn][:40]
    for k,m in list(d.items())[-4:]:
        print('coop',coop,k,names[k]); print('   ', ' '.join(f'{a.split(".")[0].replace("smsp__average_warps_issue_stalled_","stall_").replace("launch__","")[-28:]}={b}' for a,b in m.items()))
P
