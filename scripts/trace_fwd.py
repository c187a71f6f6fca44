"""Developer diagnostic: per-worker tile-start timestamps of the forward kernel -> drift between workers that stream
the same vocab split.  Usage: python scripts/trace_fwd.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, numpy as np
dev = "cuda"
trace = torch.zeros(160, 4096, 2, dtype=torch.int64, device=dev)
os.environ["VPG_DEBUG_TRACE_PTR"] = str(trace.data_ptr())
from verl_b200 import ops
T, H, V = 65536, 3584, 152064
g = torch.Generator(device=dev).manual_seed(0)
hidden = (torch.rand(T, H, device=dev, generator=g) - 0.5).bfloat16()
weight = (torch.rand(V, H, device=dev, generator=g) - 0.5).bfloat16()
labels = torch.randint(0, V, (T,), device=dev, generator=g)
for it in range(2):
    trace.zero_()
    part = ops.lce_forward_partials(hidden, weight, labels, 1.0)
    torch.cuda.synchronize()
S = part.shape[1]
tr = trace.cpu().numpy()
nw = int((tr[:, 0, 0] != 0).sum())
t0 = tr[:nw, 0, 0].min()
print("workers", nw, "splits", S, "kernel start spread us", (tr[:nw, 0, 0].max() - t0) / 1e3)
ts = (tr[:nw, :, 0] - t0) / 1e3  # us
u = tr[:nw, :, 1] >> 32
n = tr[:nw, :, 1] & 0xFFFFFFFF
valid = tr[:nw, :, 0] != 0
ntiles = valid.sum(1)
print("tiles per worker min/max", ntiles.min(), ntiles.max(), "total ms", ts[valid].max() / 1e3)
# for sampled n tiles: spread of start times among the workers that process that n tile in the same round
for rnd in (0, 3, 7, 12):
    for sp in range(S):
        starts = []
        for w in range(nw):
            uu = w + rnd * nw
            if uu % S != sp: continue
            sel = valid[w] & (u[w] == uu)
            idx = np.nonzero(sel)[0]
            if len(idx) == 0: continue
            k = idx[len(idx) // 2]
            starts.append((ts[w, k], int(n[w, k])))
        if starts:
            st = np.array([s for s, _ in starts]); nn = {x for _, x in starts}
            print(f"round {rnd} split {sp}: {len(starts)} workers, mid-unit tile {sorted(nn)[:3]}, start spread {st.max()-st.min():.1f} us (tile period ~{(ts[0, 100]-ts[0, 50]) / 50:.1f} us)")
per = (ts[:, 140] - ts[:, 10]) / 130.0
print("per-worker tile period (us), tiles 10..140 of the first unit, by physical worker id:")
print(np.array2string(per, precision=2, max_line_width=200))
per2 = (ts[:, 1900] - ts[:, 1000]) / 900.0
print("tiles 1000..1900:")
print(np.array2string(per2, precision=2, max_line_width=200))
