set -x
mkdir -p gpurun_out
N=${N:-2}
T=r2t
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29521 scripts/run_ref_lce_test.py --mode ours --tp > gpurun_out/${T}_ref_tp_test_ours_nccl_n${N}.log 2>&1; echo "ref tp test (nccl exchange) rc=$?"
VPG_TP_EXCHANGE=nvlink timeout 600 $TR --master-port 29522 scripts/run_ref_lce_test.py --mode ours --tp > gpurun_out/${T}_ref_tp_test_ours_nvlink_n${N}.log 2>&1; echo "ref tp test (nvlink exchange) rc=$?"
grep -h "PASS\|INFO\]: .*average\|Error\|error" gpurun_out/${T}_ref_tp_test_ours_*_n${N}.log | tail -40
