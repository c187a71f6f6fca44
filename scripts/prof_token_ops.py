"""One invocation of every memory-bound token op at 2^25 tokens (inputs >> L2) — the program ncu captures for the
HBM-roofline evidence of the token kernels:
    ncu --set full --clock-control none --import-source on -o gpurun_out/tok python scripts/prof_token_ops.py
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scripts"))

import torch  # noqa: E402

import bench_token_ops as b  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    L = 16384
    B = (1 << 25) // L
    d = b.make(B, L, dev)
    torch.cuda.synchronize()
    for name, fn, _ in b.ops_table(d, B, L, 8):
        torch.cuda.nvtx.range_push(name)
        fn()
        torch.cuda.synchronize()
        torch.cuda.nvtx.range_pop()
    print("done")


if __name__ == "__main__":
    main()
