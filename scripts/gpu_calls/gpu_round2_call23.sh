set -x
mkdir -p gpurun_out
timeout 300 python scripts/prof_host_overhead.py > gpurun_out/r2ag_host_overhead.txt 2>&1; echo rc=$?
