# round 2, call 34: dynamic-batch helpers after their rewrite (GPU tests of seqlen_balancing)
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -q -k "seqlen" > gpurun_out/r2ar_pytest_seqlen.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/r2ar_pytest_seqlen.log
