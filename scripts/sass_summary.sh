#!/usr/bin/env bash
# Proof of the Blackwell path in the built library: per tc_gemm_kernel instantiation, the count of tcgen05 / TMEM / TMA /
# multimem SASS mnemonics (B200_PROFILING.md: UTCHMMA = tcgen05.mma kind::f16, UTCQMMA... ; LDTM = tcgen05.ld;
# UTMALDG / UTMASTG = cp.async.bulk.tensor; UBLKCP = cp.async.bulk; UTCBAR = tcgen05.commit).
#   bash scripts/sass_summary.sh [lib] > profiles/r2_sass_summary.txt
set -euo pipefail
LIB=${1:-verl_b200/lib/libvpg_b200.so}
echo "# $(basename "$LIB"): $(cuobjdump -lelf "$LIB" | grep -c sm_100a) sm_100a cubin(s); arch list: $(cuobjdump -lelf "$LIB" | sed -E 's/.*(sm_[0-9a-z]+).*/\1/' | sort -u | tr '\n' ' ')"
cuobjdump -sass "$LIB" | awk '
  /Function :/ { fn=$3; next }
  { if (fn=="") next
    if ($0 ~ /UTCHMMA\.2CTA/) c[fn,"UTCHMMA.2CTA"]++; else if ($0 ~ /UTCHMMA/) c[fn,"UTCHMMA"]++;
    if ($0 ~ /UTCQMMA|UTCIMMA|UTCOMMA/) c[fn,"UTC?MMA(other)"]++;
    if ($0 ~ /LDTM/) c[fn,"LDTM"]++;
    if ($0 ~ /UTMALDG/) c[fn,"UTMALDG"]++;
    if ($0 ~ /UTMASTG/) c[fn,"UTMASTG"]++;
    if ($0 ~ /UBLKCP/) c[fn,"UBLKCP"]++;
    if ($0 ~ /UTCBAR/) c[fn,"UTCBAR"]++;
    if ($0 ~ /SYNCS/) c[fn,"SYNCS(mbarrier)"]++;
    if ($0 ~ /MULTIMEM|\.MMA\.RED|LDGMC|REDG\.MC|multimem/) c[fn,"multimem"]++;
    if ($0 ~ /MUFU\.EX2/) c[fn,"MUFU.EX2"]++;
    fns[fn]=1 }
  END { n=split("UTCHMMA UTCHMMA.2CTA UTC?MMA(other) LDTM UTMALDG UTMASTG UBLKCP UTCBAR SYNCS(mbarrier) multimem MUFU.EX2", k, " ");
        for (f in fns) { line=""; tot=0; for (i=1;i<=n;i++) { v=c[f,k[i]]+0; if (v>0) { line=line " " k[i] "=" v; if (k[i] != "MUFU.EX2" && k[i] != "SYNCS(mbarrier)") tot+=v } }
          if (tot>0) print f ":" line } }' | sort | c++filt | sed -E 's/void vpg::|vpg:://g'
