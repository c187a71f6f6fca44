#!/bin/bash
# usage: scripts/ncu_metrics.sh "<ENV>" tag "<python cmd>" count  -> per-launch metrics table
M=gpu__time_duration.sum,dram__bytes_read.sum,lts__t_sector_hit_rate.pct,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__cycles_elapsed.avg.per_second
env $1 $3 > gpurun_out/m_$2.plain 2>&1 && \
env $1 ncu --metrics $M --clock-control none -k regex:tc_gemm -c $4 --csv --log-file gpurun_out/m_$2.csv $3 > /dev/null 2>&1
python - "$2" <<'PY'
import csv,sys
tag=sys.argv[1]
rows=[r for r in csv.reader(open(f'gpurun_out/m_{tag}.csv')) if len(r)>10]
h=rows[0]; n=h.index('Metric Name'); v=h.index('Metric Value'); i=h.index('ID'); kn=h.index('Kernel Name')
d={}; names={}
for r in rows[1:]:
    d.setdefault(r[i],{})[r[n]]=r[v]; names[r[i]]=r[kn][:34]
for k,m in d.items():
    print(tag, k, names[k], ' '.join(f'{a.split(".")[0][-22:]}={b}' for a,b in m.items()))
PY
