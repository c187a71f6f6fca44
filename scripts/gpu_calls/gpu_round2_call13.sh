set -x
mkdir -p gpurun_out
T=r2u
run() { tag=$1; shift; env "$@" timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-parity > gpurun_out/${T}_bench_$tag.json 2> gpurun_out/${T}_bench_$tag.err; echo "$tag rc=$?"; python scripts/show_bench.py gpurun_out/${T}_bench_$tag.json | head -2; }
run base A=1
run stcs VPG_DS_STCS=1
run stcs_hintA VPG_DS_STCS=1 VPG_HINT_A_DW=1 VPG_HINT_A_DH=1
run hintB_last VPG_HINT_B_DW=2 VPG_HINT_B_DH=2
run base2 A=1
