"""Plain tcgen05 GEMM launches over operand layouts and K lengths (for ncu metric sweeps of the mainloop)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from verl_b200 import ops
dev = "cuda"
M, N = 37888, 4096   # 148 x 256 rows, 16 n tiles
for K in (3584, 14336):
    A = (torch.rand(M, K, device=dev) - 0.5).bfloat16()
    B = (torch.rand(N, K, device=dev) - 0.5).bfloat16()
    At, Bt = A.t().contiguous(), B.t().contiguous()
    for a_k, b_k in ((True, True), (True, False), (False, False)):
        C = ops.gemm_bf16(A if a_k else At, B if b_k else Bt, a_k, b_k)
    torch.cuda.synchronize()
    del A, B, At, Bt
print("ok", float(C.float().abs().mean()))
