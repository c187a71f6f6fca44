#!/bin/bash
# usage: scripts/ncu_sweep.sh "<ENV assignments>" tag   -> gpurun_out/sweep_<tag>.csv (fwd kernel metrics)
M=dram__bytes_read.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__cycles_elapsed.avg.per_second
env $1 python scripts/prof_fwd.py > gpurun_out/sweep_$2.plain 2>&1 && \
env $1 ncu --metrics $M --clock-control none -k regex:tc_gemm -c 2 --csv --log-file gpurun_out/sweep_$2.csv python scripts/prof_fwd.py > /dev/null 2>&1
python - "$2" <<'PY'
import csv,sys
tag=sys.argv[1]
rows=[r for r in csv.reader(open(f'gpurun_out/sweep_{tag}.csv')) if len(r)>10]
h=rows[0]; n=h.index('Metric Name'); v=h.index('Metric Value'); i=h.index('ID')
d={}
for r in rows[1:]:
    d.setdefault(r[i],{})[r[n]]=r[v]
for k,m in d.items():
    print(tag, k, ' '.join(f'{a.split(".")[0][-28:]}={b}' for a,b in m.items()))
PY
