"""Developer probe run on the GPU box: exercises the tcgen05 kernels step by step, each step in its own
subprocess with a timeout so that a hang or fault in one step does not hide the others.

    python scripts/gpu_probe.py            # all steps
    python scripts/gpu_probe.py gemm fwd   # selected steps
"""

from __future__ import annotations

import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _ref_lce(hidden, weight, labels, temperature):
    import torch

    logits = (hidden.float() @ weight.float().t()) / temperature
    lse = torch.logsumexp(logits, dim=-1)
    p = torch.softmax(logits, dim=-1)
    ent = lse - (p * logits).sum(-1)
    lp = logits.gather(-1, labels[:, None]).squeeze(-1) - lse
    return lp, ent


def step_gemm():
    import torch

    from verl_b200 import _lib

    lib = _lib.load()
    torch.manual_seed(0)
    dev = "cuda"
    ok = True
    for (M, N, K) in [(128, 256, 64), (300, 520, 192), (1000, 3584, 1024)]:
        for a_k in (1, 0):
            for b_k in (1, 0):
                A = (torch.rand(M, K, device=dev) - 0.5).bfloat16()
                B = (torch.rand(N, K, device=dev) - 0.5).bfloat16()
                ref = A.float() @ B.float().t()
                # pad leading dims to multiples of 8 for TMA
                def store(x, kmajor):
                    if kmajor:
                        return x.contiguous(), x.shape[1]
                    xt = x.t().contiguous()  # [K, rows]
                    rows = xt.shape[1]
                    ld = (rows + 7) // 8 * 8
                    buf = torch.zeros(xt.shape[0], ld, device=dev, dtype=x.dtype)
                    buf[:, :rows] = xt
                    return buf, ld
                As, lda = store(A, a_k)
                Bs, ldb = store(B, b_k)
                ldc = (N + 7) // 8 * 8
                Cbuf = torch.zeros(M, ldc, device=dev, dtype=torch.float32)
                _lib.check(lib.vpg_gemm_bf16(As.data_ptr(), Bs.data_ptr(), M, N, K, lda, ldb, a_k, b_k,
                                             Cbuf.data_ptr(), ldc, 1, None, 0, _lib.stream_ptr()))
                torch.cuda.synchronize()
                err = (Cbuf[:, :N] - ref).abs().max().item()
                scale = ref.abs().max().item()
                good = err <= 1e-3 * max(scale, 1.0)
                ok &= good
                print(f"gemm M={M} N={N} K={K} a_k={a_k} b_k={b_k}: max_err={err:.3e} scale={scale:.2f} {'OK' if good else 'FAIL'}", flush=True)
    return ok


def step_fwd():
    import torch

    from verl_b200 import _lib

    lib = _lib.load()
    dev = "cuda"
    ok = True
    for (T, H, V, temp, splits) in [(128, 64, 256, 1.0, 1), (300, 128, 1000, 1.5, 2), (1937, 3584, 152064 // 8, 1.5, 0),
                                    (2169, 896, 151936, 1.5, 0)]:
        torch.manual_seed(1)
        hidden = (torch.rand(T, H, device=dev) - 0.5).bfloat16()
        weight = (torch.rand(V, H, device=dev) - 0.5).bfloat16()
        labels = torch.randint(0, V, (T,), device=dev)
        S = splits or lib.vpg_lce_num_splits(T, H, V, 0)
        part = torch.empty(4, S, T, device=dev, dtype=torch.float32)
        outs = [torch.empty(T, device=dev, dtype=torch.float32) for _ in range(4)]
        _lib.check(lib.vpg_lce_fwd_partials(hidden.data_ptr(), weight.data_ptr(), labels.data_ptr(), T, H, V, temp, 0,
                                            S, part.data_ptr(), _lib.stream_ptr()))
        _lib.check(lib.vpg_lce_finalize(part.data_ptr(), S, T, temp, *[o.data_ptr() for o in outs],
                                        _lib.stream_ptr()))
        torch.cuda.synchronize()
        lp_ref, ent_ref = _ref_lce(hidden, weight, labels, temp)
        e1 = (outs[0] - lp_ref).abs().max().item()
        e2 = (outs[1] - ent_ref).abs().max().item()
        good = e1 < 2e-3 and e2 < 2e-3
        ok &= good
        print(f"fwd T={T} H={H} V={V} S={S}: |dlogprob|={e1:.3e} |dentropy|={e2:.3e} {'OK' if good else 'FAIL'}", flush=True)
    return ok


def step_bwd():
    import torch

    from verl_b200 import _lib

    lib = _lib.load()
    dev = "cuda"
    ok = True
    for (T, H, V, temp, chunk) in [(128, 64, 256, 1.0, 0), (300, 128, 1000, 1.5, 512), (1937, 1024, 19008, 1.5, 0)]:
        torch.manual_seed(2)
        hidden = (torch.rand(T, H, device=dev) - 0.5).bfloat16().requires_grad_(True)
        weight = (torch.rand(V, H, device=dev) - 0.5).bfloat16().requires_grad_(True)
        labels = torch.randint(0, V, (T,), device=dev)
        g_lp = (torch.rand(T, device=dev) * 2 - 1)
        g_ent = (torch.rand(T, device=dev) - 0.5)
        hf = hidden.detach().float().requires_grad_(True)
        wf = weight.detach().float().requires_grad_(True)
        lp_ref, ent_ref = _ref_lce(hf, wf, labels, temp)
        (lp_ref * g_lp + ent_ref * g_ent).sum().backward()
        S = lib.vpg_lce_num_splits(T, H, V, 0)
        part = torch.empty(4, S, T, device=dev, dtype=torch.float32)
        outs = [torch.empty(T, device=dev, dtype=torch.float32) for _ in range(4)]
        st = _lib.stream_ptr()
        _lib.check(lib.vpg_lce_fwd_partials(hidden.data_ptr(), weight.data_ptr(), labels.data_ptr(), T, H, V, temp, 0,
                                            S, part.data_ptr(), st))
        _lib.check(lib.vpg_lce_finalize(part.data_ptr(), S, T, temp, *[o.data_ptr() for o in outs], st))
        ws_bytes = lib.vpg_lce_bwd_workspace_bytes(T, H, V, chunk)
        ws = torch.empty(ws_bytes, device=dev, dtype=torch.uint8)
        dh = torch.empty_like(hidden)
        dw = torch.empty_like(weight)
        _lib.check(lib.vpg_lce_bwd(hidden.data_ptr(), weight.data_ptr(), labels.data_ptr(), outs[2].data_ptr(),
                                   outs[3].data_ptr(), g_lp.data_ptr(), g_ent.data_ptr(), T, H, V, temp, 0, chunk,
                                   dh.data_ptr(), dw.data_ptr(), ws.data_ptr(), ws_bytes, st))
        torch.cuda.synchronize()
        e1 = (dh.float() - hf.grad).abs().max().item()
        e2 = (dw.float() - wf.grad).abs().max().item()
        s1, s2 = hf.grad.abs().max().item(), wf.grad.abs().max().item()
        good = e1 <= 2e-2 + 4e-2 * s1 and e2 <= 2e-2 + 4e-2 * s2
        ok &= good
        print(f"bwd T={T} H={H} V={V} chunk={chunk}: |d_hidden err|={e1:.3e} (max {s1:.2e}) |d_weight err|={e2:.3e} (max {s2:.2e}) {'OK' if good else 'FAIL'}", flush=True)
    return ok


def step_time():
    import torch

    from verl_b200 import _lib

    lib = _lib.load()
    dev = "cuda"
    T, H, V = 65536, 3584, 152064
    torch.manual_seed(3)
    hidden = (torch.rand(T, H, device=dev) - 0.5).bfloat16()
    weight = (torch.rand(V, H, device=dev) - 0.5).bfloat16()
    labels = torch.randint(0, V, (T,), device=dev)
    st = _lib.stream_ptr()
    for S in [0, 2, 4, 8, 16]:
        S_ = S or lib.vpg_lce_num_splits(T, H, V, 0)
        part = torch.empty(4, S_, T, device=dev, dtype=torch.float32)
        outs = [torch.empty(T, device=dev, dtype=torch.float32) for _ in range(4)]

        def run():
            _lib.check(lib.vpg_lce_fwd_partials(hidden.data_ptr(), weight.data_ptr(), labels.data_ptr(), T, H, V, 1.0,
                                                0, S_, part.data_ptr(), st))
            _lib.check(lib.vpg_lce_finalize(part.data_ptr(), S_, T, 1.0, *[o.data_ptr() for o in outs], st))

        for _ in range(2):
            run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        n = 3
        for _ in range(n):
            run()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        tf = 2.0 * T * H * V / (ms * 1e-3) / 1e12
        print(f"time fwd C2 S={S_}: {ms:.2f} ms  {tf:.1f} TFLOP/s  {T / (ms * 1e-3) / 1e6:.3f} Mtok/s", flush=True)
    # backward
    g_lp = torch.rand(T, device=dev) * 2 - 1
    g_ent = torch.rand(T, device=dev) - 0.5
    ws_bytes = lib.vpg_lce_bwd_workspace_bytes(T, H, V, 0)
    ws = torch.empty(ws_bytes, device=dev, dtype=torch.uint8)
    dh = torch.empty_like(hidden)
    dw = torch.empty_like(weight)

    def runb():
        _lib.check(lib.vpg_lce_bwd(hidden.data_ptr(), weight.data_ptr(), labels.data_ptr(), outs[2].data_ptr(),
                                   outs[3].data_ptr(), g_lp.data_ptr(), g_ent.data_ptr(), T, H, V, 1.0, 0, 0,
                                   dh.data_ptr(), dw.data_ptr(), ws.data_ptr(), ws_bytes, st))

    runb()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    runb()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print(f"time bwd C2: {ms:.2f} ms  useful {4.0 * T * H * V / (ms * 1e-3) / 1e12:.1f} TFLOP/s executed {6.0 * T * H * V / (ms * 1e-3) / 1e12:.1f} TFLOP/s", flush=True)
    # cuBLAS comparator for context
    a = hidden[:8192]
    torch.matmul(a, weight.t())
    torch.cuda.synchronize()
    e0.record()
    for _ in range(3):
        torch.matmul(a, weight.t())
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    print(f"cuBLAS bf16 [8192x{H}]x[{H}x{V}]: {ms:.2f} ms {2.0 * 8192 * H * V / (ms * 1e-3) / 1e12:.1f} TFLOP/s", flush=True)
    return True


STEPS = {"gemm": step_gemm, "fwd": step_fwd, "bwd": step_bwd, "time": step_time}

if __name__ == "__main__":
    if len(sys.argv) >= 3 and sys.argv[1] == "--child":
        ok = STEPS[sys.argv[2]]()
        sys.exit(0 if ok else 1)
    names = sys.argv[1:] or list(STEPS)
    rc = 0
    for name in names:
        t0 = time.time()
        try:
            r = subprocess.run([sys.executable, __file__, "--child", name], timeout=int(os.environ.get("PROBE_TIMEOUT", "240")))
            code = r.returncode
        except subprocess.TimeoutExpired:
            code = -999
        print(f"== step {name}: exit {code} in {time.time() - t0:.1f}s", flush=True)
        rc |= 1 if code else 0
    sys.exit(rc)
