"""A/B of the narrow-row pad_input launch geometry (VPG_PAD_NT / VPG_PAD_TILE_KB / VPG_PAD_BPS are read once per process,
so every setting runs in its own subprocess).  Prints one JSON line per setting."""
import json, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import json, os, sys, torch
sys.path.insert(0, %r)
from verl_b200 import ops
sys.path.insert(0, os.path.join(%r, "scripts"))
import bench_token_ops as bto
dev = torch.device("cuda:0")
L = 16384; B = (1 << 25) // L
d = bto.make(B, L, dev)
idx, _, _, _ = ops.unpad_indices(d["mask"])
nnz = idx.numel(); n = B * L
vals = d["lp"].reshape(-1, 1)[:nnz].contiguous()
ms = bto.timeit(lambda: ops.scatter_rows(vals, idx, n), iters=20, warmup=5)
ref = torch.zeros(n, 1, device=dev); ref[idx] = vals
ok = bool(torch.equal(ops.scatter_rows(vals, idx, n), ref))
nb = nnz * 12 + n * 4
print(json.dumps({"env": {k: v for k, v in os.environ.items() if k.startswith("VPG_PAD")}, "ms": round(ms, 4), "gbs": round(nb / ms / 1e6, 1), "ok": ok}))
''' % (ROOT, ROOT)
for nt in (128, 256, 512, 1024):
    for kb in (8, 16, 32):
        for bps in (2, 4, 6, 8, 10, 12, 16):
            if nt * bps > 2048 or kb * bps > 200 or (nt >= 512 and kb < 32) or nt * bps < 1024:
                continue
            env = dict(os.environ, VPG_PAD_NT=str(nt), VPG_PAD_TILE_KB=str(kb), VPG_PAD_BPS=str(bps))
            r = subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True)
            print(r.stdout.strip().splitlines()[-1] if r.stdout.strip() else "ERR " + r.stderr[-300:], flush=True)
