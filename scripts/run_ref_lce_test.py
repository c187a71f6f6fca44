"""Run the reference's OWN lm-head kernel test script (tests/utils/test_linear_cross_entropy.py, snapshot under
oracle/_ref) on this GPU, unmodified, in one of two modes:

  --mode reference   every import resolves to the reference snapshot (its Triton kernel is the "Kernel implementation"):
                     the reference's own latencies / peak memory on this B200;
  --mode ours        `verl.utils.kernel.linear_cross_entropy` (and `.kernels`) resolve to verl_b200, everything else stays
                     the reference: the reference's assertions (logprobs atol 1e-3, entropy atol 5e-3, grads atol 2e-2 /
                     rtol 4e-2 against three PyTorch implementations, 5 shapes) judge the tcgen05 kernels.

    python scripts/run_ref_lce_test.py --mode ours
    torchrun --nproc-per-node 2 --master-addr 127.0.0.1 scripts/run_ref_lce_test.py --mode ours --tp

--tp runs the reference's TENSOR-PARALLEL test script (tests/utils/test_linear_cross_entropy_tp.py) instead: every rank
holds its own vocab shard, the sharded op is compared with PyTorch on the all-gathered weight (logprobs / entropy atol 1e-1
rtol 1e-2 as the script states them, grads atol 2e-2 rtol 4e-2), d_hidden all-reduced by the script itself.  With
VPG_TP_EXCHANGE=nvlink the forward's exchange is the NVLink-fused one (symmetric buffers created collectively on first use).
"""
import argparse
import os
import runpy
import sys

# The script compiles FusedLinearForPPO with torch.compile; torch 2.11's dynamo trips an internal assertion on that
# autograd.Function (higher_order_ops.overwrite_tensor_vt_proxy) — an incompatibility between the reference's test and this
# image's torch, in BOTH modes.  Dynamo is therefore disabled (torch.compile becomes a no-op, the script itself stays
# unmodified): the PyTorch comparators run eagerly, which changes their speed, not their results.
os.environ.setdefault("TORCHDYNAMO_DISABLE", "1")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", choices=["reference", "ours"], required=True)
    ap.add_argument("--tp", action="store_true", help="the reference's tensor-parallel test script (run under torchrun)")
    args = ap.parse_args()
    if args.tp:
        import torch

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    from oracle import ref_import

    ref_import.load()  # verl, verl.utils, verl.utils.torch_functional, verl.utils.experimental.* from the snapshot
    if args.mode == "reference":
        ref_import.load_kernel()
    else:
        import types

        from verl_b200.utils.kernel import kernels as k
        from verl_b200.utils.kernel import linear_cross_entropy as lce

        pkg = types.ModuleType("verl.utils.kernel")
        pkg.__path__ = []
        sys.modules["verl.utils.kernel"] = pkg
        sys.modules["verl.utils.kernel.kernels"] = k
        sys.modules["verl.utils.kernel.linear_cross_entropy"] = lce
        pkg.kernels, pkg.linear_cross_entropy = k, lce
        sys.modules["verl.utils"].kernel = pkg
    script = os.path.join(ref_import.REFERENCE_ROOT, "tests", "utils",
                          "test_linear_cross_entropy_tp.py" if args.tp else "test_linear_cross_entropy.py")
    print(f"[run_ref_lce_test] mode={args.mode} script={script}", flush=True)
    runpy.run_path(script, run_name="__main__")
    print(f"[run_ref_lce_test] mode={args.mode}: all cases passed", flush=True)


if __name__ == "__main__":
    main()
