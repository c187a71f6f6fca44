import torch
dev="cuda"
M,N=37888,4096
for K in (3584,14336):
    A=(torch.rand(M,K,device=dev)-0.5).bfloat16(); B=(torch.rand(N,K,device=dev)-0.5).bfloat16()
    for _ in range(2): C=A@B.t()
    torch.cuda.synchronize()
print("ok")
