"""Device time of vpg_unpad_indices (no host sync) on a 2^25-element int64 mask (S = 8192), CUDA events."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from verl_b200 import ops
from verl_b200._lib import load
from verl_b200.ops import ptr, stream_ptr, mask_arg, check

B, S = 4096, 8192
dev = torch.device("cuda")
g = torch.Generator(device="cuda").manual_seed(0)
lens = torch.randint(S // 4, S + 1, (B,), device="cuda", generator=g)
mask = (torch.arange(S, device="cuda")[None] < lens[:, None]).long()
lib = load()
n = B * S
indices = torch.empty(n, dtype=torch.int64, device=dev)
seqlens = torch.empty(B, dtype=torch.int32, device=dev)
cu = torch.empty(B + 1, dtype=torch.int32, device=dev)
totals = torch.zeros(2, dtype=torch.int32, device=dev)
ws = torch.empty(int(lib.vpg_unpad_workspace_bytes(n)) // 8 + 2, dtype=torch.int64, device=dev)
m, mdt = mask_arg(mask)


def call():
    check(lib.vpg_unpad_indices(ptr(m), mdt, B, S, ptr(indices), ptr(seqlens), ptr(cu), ptr(totals), ptr(ws),
                                ws.numel() * 8, stream_ptr(dev)))


for _ in range(3):
    call()
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
ev[0].record()
for _ in range(20):
    call()
ev[1].record()
torch.cuda.synchronize()
us = ev[0].elapsed_time(ev[1]) / 20 * 1e3
nnz = int(totals[0])
print(f"unpad_indices device us/call {us:.1f}  nnz {nnz}"
      f"  ({(n * 8 + nnz * 8) / us / 1e3:.0f} GB/s algorithmic)")
if True:
    assert torch.equal(indices[:nnz], mask.flatten().nonzero().flatten())
    assert torch.equal(seqlens.long(), mask.sum(-1))
