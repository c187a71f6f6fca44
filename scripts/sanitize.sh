#!/usr/bin/env bash
# compute-sanitizer over the small-shape GPU parity tests of the lm-head / rmpad / token kernels, for both
# tile variants (VPG_CTA_GROUP=2 default, =1).  Writes one summary per (tool, variant) under gpurun_out/.
#   bash scripts/sanitize.sh [memcheck racecheck synccheck]
# Sanitised runs are 10-100x slower: only the golden / small-shape tests are selected.
set -uo pipefail
OUT=${OUT:-gpurun_out}
mkdir -p "$OUT"
TOOLS=${*:-memcheck racecheck synccheck}
SEL='golden or known_answers or all_masked or unaligned or reductions_and_3d or split_invariance or multi_turn or autograd_of_index or gemm_all_majors'
FILES="tests/test_lce_gpu.py tests/test_rmpad_gpu.py tests/test_token_ops_gpu.py"
for cg in 2 1; do
  for tool in $TOOLS; do
    log="$OUT/r2_sanitizer_${tool}_cta${cg}.log"
    extra=""
    [ "$tool" = memcheck ] && extra="--leak-check no"
    VPG_CTA_GROUP=$cg timeout ${SAN_TIMEOUT:-900} compute-sanitizer --tool "$tool" $extra --print-limit 20 \
        --target-processes application-only \
        python -m pytest $FILES -m gpu -q -x -k "$SEL" -p no:cacheprovider > "$log" 2>&1
    rc=$?
    {
      echo "== $tool, VPG_CTA_GROUP=$cg: exit code $rc"
      grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|error" "$log" | tail -6
    } | tee -a "$OUT/r2_sanitizer_summary.txt"
  done
done
