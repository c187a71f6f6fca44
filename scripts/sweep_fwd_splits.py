"""Forward lm-head kernel time against the number of vocab splits per token tile (units = token tile x vocab range):
more splits = fewer token tiles in flight (smaller A footprint in L2) but more W streams.  Prints one JSON line per
(shape, splits); the library's own choice is marked."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from verl_b200 import ops

dev = torch.device("cuda:0")
shapes = {"c2": (65536, 3584, 152064), "c4": (65536, 8192, 128256), "c5s": (262144, 5120, 19008), "c4s8": (65536, 8192, 16032),
          "c2s2": (65536, 3584, 76032), "c2s4": (65536, 3584, 38016), "c2s8": (65536, 3584, 19008),
          "c4s2": (65536, 8192, 64128), "c4s4": (65536, 8192, 32064), "c1": (20480, 896, 151936)}
ONLY = [int(x) for x in os.environ.get("SWEEP_SPLITS", "").split(",") if x]
for name in (sys.argv[1:] or list(shapes)):
    T, H, V = shapes[name]
    g = torch.Generator(device=dev).manual_seed(0)
    hidden = (torch.rand(T, H, device=dev, generator=g) - 0.5).bfloat16()
    weight = (torch.rand(V, H, device=dev, generator=g) - 0.5).bfloat16()
    labels = torch.randint(0, V, (T,), device=dev, generator=g)
    auto = ops.lce_num_splits(T, H, V)
    n_tiles = -(-V // 256)
    cands = sorted(({auto} | set(ONLY or [1, 2, 3, 4, 6, 8, 12, 16, 24, 32])) & set(range(1, n_tiles + 1)))
    for S in cands:
        tps = -(-n_tiles // S)
        if -(-n_tiles // tps) != S:
            continue
        for _ in range(3):
            ops.lce_forward_partials(hidden, weight, labels, 1.0, num_splits=S)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = max(3, int(1500 / (2e-9 * T * H * V / 1.3e3 + 1)))
        e0.record()
        for _ in range(reps):
            ops.lce_forward_partials(hidden, weight, labels, 1.0, num_splits=S)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        print(json.dumps({"shape": name, "T": T, "H": H, "V": V, "splits": S, "auto": S == auto, "ms": round(ms, 3),
                          "tflops": round(2.0 * T * H * V / ms / 1e9, 1)}), flush=True)
    del hidden, weight
