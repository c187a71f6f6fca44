# round 2, call 35: the full GPU suite and smoke at HEAD (3072 MiB default dlogits budget, rewritten seqlen helpers)
set -x
mkdir -p gpurun_out
T=r2as
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${T}_pytest_gpu.log; tail -n 4 gpurun_out/${T}_pytest_gpu.log
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/${T}_smoke.log
