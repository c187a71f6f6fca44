#!/usr/bin/env python
"""Benchmark of the per-token policy-gradient hot path (BASELINE.json metric: fused lm-head logprob+entropy tokens/s
at Qwen2.5-7B shape, 1-8 B200).

One "step" = one pass of the hot path over one rmpad micro-batch of synthetic hidden states:
    fused lm-head forward (log-prob + entropy, logits never materialised)
    -> fused PPO actor loss (clip / dual-clip, entropy bonus, low-var KL, token-mean) forward + closed-form gradient
    -> fused lm-head backward (recompute tiles -> d_hidden, d_weight)
through the public API (linear_cross_entropy autograd op + core_algos.fused_actor_loss).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--tokens T]

N = 1: BASELINE configs[1] (H=3584, V=152064, T=65536 rmpad tokens).  N > 1 (torchrun): the same micro-batch with the
lm-head weight vocab-parallel over N ranks (V/N rows each), one all-gather of the packed per-token partials in the
forward and one all-reduce of d_hidden in the backward -> strong scaling.
--impl reference: the reference algorithm's CPU path (oracle port, all host threads) on a bounded token sample.
Prints ONE JSON line on rank 0.
"""

from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

HIDDEN, VOCAB, TOKENS = 3584, 152064, 65536  # Qwen2.5-7B lm-head, rmpad 64k tokens (BASELINE.json configs[1])
METRIC = "fused lm-head logprob+entropy fwd+bwd tokens/s (Qwen2.5-7B shape)"
LOSS_CFG = dict(cliprange=0.2, cliprange_low=0.2, cliprange_high=0.28, clip_ratio_c=3.0, loss_agg_mode="token-mean",
                entropy_coeff=1e-3, kl_loss_coef=0.01, kl_loss_type="low_var_kl")


def workload_config(T, world, exchange="nvlink"):
    """`config` of the JSON line — identical for both arms (the reference arm times a bounded sample of it)."""
    red = ("NCCL all-reduce of d_hidden (on a second stream under the d_weight GEMM where the cost model says the "
           "hidden all-reduce pays for the SMs that GEMM spares; see the line's tp_reduce)")
    how = ("partials pushed into every rank's gathered buffer by the forward GEMM epilogue (NVSwitch multimem.st, no "
           "separate all-gather) + " + red
           if exchange == "nvlink" else "NCCL all-gather of [4,S,T] partials + " + red)
    return {
        "workload": f"Qwen2.5-7B lm-head (H={HIDDEN}, V={VOCAB}), rmpad T={T} tokens: fused logprob+entropy fwd -> "
                    "PPO clip/dual-clip + entropy bonus + low_var_kl token-mean loss -> fused bwd (d_hidden, d_weight)",
        "parallelism": "single GPU" if world == 1 else f"vocab-parallel lm-head over {world} GPUs (V/{world} rows each), " + how,
        "l2": "inputs exceed L2 (hidden 470 MB + weight shard >= 136 MB per step)",
        "loss": LOSS_CFG,
    }


def _traffic(T, H, V):
    """DRAM bytes per forward launch from the committed `ncu --set full` capture (None if the shape differs)."""
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            t = json.load(f)["tc_gemm_fwd"]
        if (t["T"], t["H"], t["V"]) == (T, H, V):
            return t["dram_read_bytes"] + t["dram_write_bytes"]
    except Exception:
        pass
    return None


def _peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return dict(burst=float(p["bf16_tflops"]), sustained=float(p.get("bf16_tflops_sustained", p["bf16_tflops"])),
                    hbm=float(p["hbm_gbs"]), source="measured (MEASURED_PEAKS.json)")
    except Exception:
        return dict(burst=1590.0, sustained=1400.0, hbm=6650.0, source="fallback (B200_PROFILING.md)")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs"""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = sorted(float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit())
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 7 for n, v in zip(names, r[3:7]) if v.lower() == "active"})
        pw = [float(r[2]) for r in self.rows if len(r) > 2 and r[2].replace(".", "").isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": reasons}


# ----------------------------------------------------------------------------------------------- CPU reference arm
_CPU_INPUTS: dict = {}


def cpu_reference_sample(sample_tokens: int, threads: int, repeats: int = 1):
    """The reference algorithm on the host cores: chunked (512) lm-head log-prob/entropy forward + hand-written
    backward (verl/utils/experimental/torch_functional.py:20-72) and the actor loss composition
    (verl/trainer/ppo/core_algos.py + dp_actor.py:422-469), via the oracle port.  Returns tokens/s."""
    import torch

    from oracle import port

    torch.set_num_threads(threads)
    if _CPU_INPUTS.get("n") != sample_tokens:  # synthetic inputs are built once, outside every timed region
        g = torch.Generator().manual_seed(233376)
        _CPU_INPUTS.update(
            n=sample_tokens,
            hidden=(torch.rand(sample_tokens, HIDDEN, generator=g) - 0.5).bfloat16(),
            w32=(torch.rand(VOCAB, HIDDEN, generator=g) - 0.5).bfloat16().float(),  # parameter cast: not part of a step
            labels=torch.randint(0, VOCAB, (sample_tokens,), generator=g),
            old=-torch.rand(1, sample_tokens, generator=g) * 3,
            adv=torch.randn(1, sample_tokens, generator=g),
            mask=(torch.rand(1, sample_tokens, generator=g) < 0.9).long(),
        )
    hidden, w32, labels = _CPU_INPUTS["hidden"], _CPU_INPUTS["w32"], _CPU_INPUTS["labels"]
    old, adv, mask = _CPU_INPUTS["old"], _CPU_INPUTS["adv"], _CPU_INPUTS["mask"]
    best = float("inf")
    for _ in range(repeats):
        t0 = time.perf_counter()
        lp, ent = port.lce_forward(hidden, w32, labels, 1.0)
        lp_ = lp.view(1, -1).clone().requires_grad_(True)
        ent_ = ent.view(1, -1).clone().requires_grad_(True)
        loss, _ = port.actor_loss(old, lp_, adv, mask, cliprange_low=0.2, cliprange_high=0.28, clip_ratio_c=3.0,
                                  entropy=ent_, entropy_coeff=1e-3, ref_log_prob=old, kl_loss_coef=0.01)
        loss.backward()
        port.lce_backward(hidden, w32, labels, lp_.grad.view(-1), ent_.grad.view(-1), 1.0)
        best = min(best, time.perf_counter() - t0)
    return sample_tokens / best, best


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    sample = args.cpu_tokens or 2048
    times = []
    for i in range(args.warmup + args.steps):
        tps, dt = cpu_reference_sample(sample, threads)
        if i >= args.warmup:
            times.append(dt)
    ms = 1e3 * sum(times) / len(times)
    value = sample / (ms * 1e-3)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "tokens/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32 (bf16 inputs upcast, CPU)", "data": "synthetic",
        "config": workload_config(args.tokens, max(args.gpus, 1)),
        "cpu_baseline": {"value": value, "unit": "tokens/s", "cores": threads, "kind": "port",
                         "sample": f"{sample}-token sample of the workload per step at full H and V, oracle port of the reference's "
                                   "chunk-512 CPU path (tokens/s is linear in T)"},
        "e2e": {"value": value, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------- our arm
def run_ours(args):
    import torch
    import torch.distributed as dist

    from verl_b200 import _lib
    from verl_b200.trainer.ppo import core_algos
    from verl_b200.utils.kernel.linear_cross_entropy import linear_cross_entropy

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: verl_b200 has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    group = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
        group = dist.group.WORLD
    _lib.load()
    exchange = "nccl"
    reduce_how = "nccl"
    if world > 1:
        from verl_b200.utils.kernel import kernels, peer

        want = args.tp_exchange
        ok = 0
        if want in ("auto", "nvlink"):
            try:  # collective pre-flight: symmetric allocation + NVSwitch multicast binding
                ok = int(peer.workspace(group, dev).has_multicast)
            except Exception as e:  # noqa: BLE001
                print(f"[bench] rank {rank}: NVLink peer workspace unavailable ({type(e).__name__}: {e}); using NCCL",
                      file=sys.stderr, flush=True)
        flag = torch.tensor([ok], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        exchange = "nvlink" if int(flag.item()) == 1 else "nccl"
        kernels.set_vocab_parallel_exchange(exchange)
        kernels.set_reduce_d_hidden(True, mode=args.tp_reduce)  # the op returns the summed d_hidden
        reduce_how = args.tp_reduce
        if reduce_how == "auto":
            reduce_how = "overlap" if kernels._overlap_plan(args.tokens, HIDDEN, VOCAB // world, world, dev)[0] else "nccl"

    T, H, V = args.tokens, HIDDEN, VOCAB
    assert V % world == 0
    v_local = V // world
    gen = torch.Generator(device=dev).manual_seed(233376)  # same on all ranks: hidden / labels are replicated
    # ---- synthetic inputs: pinned host copies for the end-to-end leg, device copies for `value`
    hidden_d = (torch.rand(T, H, device=dev, generator=gen) - 0.5).bfloat16()
    labels_d = torch.randint(0, V, (T,), device=dev, generator=gen)
    old_d = -torch.rand(1, T, device=dev, generator=gen) * 3
    ref_d = old_d + 0.05 * torch.randn(1, T, device=dev, generator=gen)
    adv_d = torch.randn(1, T, device=dev, generator=gen)
    mask_d = (torch.rand(1, T, device=dev, generator=gen) < 0.9).long()
    wgen = torch.Generator(device=dev).manual_seed(1000 + rank)
    weight = (torch.rand(v_local, H, device=dev, generator=wgen) - 0.5).bfloat16().requires_grad_(True)
    host = {k: v.cpu().pin_memory() for k, v in dict(hidden=hidden_d, labels=labels_d, old=old_d, ref=ref_d, adv=adv_d,
                                                     mask=mask_d).items()}
    # end-to-end leg: two staging slots filled by a copy stream, so the H2D copy of step i+1 runs under the compute of
    # step i (every step's copy is inside the timed region).  With N ranks `hidden` is replicated: each rank copies
    # only its T/N rows over PCIe and the ranks all-gather the rest over NVLink.
    assert T % world == 0
    rows = T // world
    staging = [{k: torch.empty_like(v, device=dev) for k, v in host.items()} for _ in range(2)]
    # N > 1: the replicated `hidden` is assembled WITHOUT SM work: its two staging slots are symmetric (peer-mapped)
    # allocations, each rank copies its T/N rows from the host and pulls the other ranks' rows with copy-engine
    # peer-to-peer DMA over NVLink on the copy stream, between two signal-pad barriers.  An NCCL all-gather here would
    # put ~16 CTAs next to a persistent 1-CTA-per-SM GEMM whose static schedule then waits for the delayed workers.
    gather = "none"
    peer_views = [None, None]
    sym_handles = [None, None]
    if world > 1:
        gather = args.e2e_gather
        if gather == "ce":
            ok = 1
            try:
                import torch.distributed._symmetric_memory as symm_mem

                try:
                    import warnings

                    with warnings.catch_warnings():  # deprecated no-op on newer torch
                        warnings.simplefilter("ignore")
                        symm_mem.enable_symm_mem_for_group(group.group_name)
                except Exception:  # noqa: BLE001  (newer torch: implicit)
                    pass
                for s_ in range(2):
                    buf = symm_mem.empty(T * H, dtype=torch.bfloat16, device=dev)
                    hdl = symm_mem.rendezvous(buf, group=group.group_name)
                    staging[s_]["hidden"] = buf.view(T, H)
                    peer_views[s_] = [hdl.get_buffer(r, (T, H), torch.bfloat16) for r in range(world)]
                    sym_handles[s_] = hdl
            except Exception as e:  # noqa: BLE001
                ok = 0
                print(f"[bench] rank {rank}: symmetric staging unavailable ({type(e).__name__}: {e}); NCCL all-gather",
                      file=sys.stderr, flush=True)
            flag = torch.tensor([ok], device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            if int(flag.item()) != 1:
                gather = "nccl"
                for s_ in range(2):
                    staging[s_]["hidden"] = torch.empty_like(host["hidden"], device=dev)
    copy_stream = torch.cuda.Stream(device=dev)
    ready = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]
    small_bytes = sum(v.numel() * v.element_size() for k, v in host.items() if k != "hidden")
    h2d_bytes = host["hidden"].numel() * host["hidden"].element_size() + world * small_bytes  # whole job, per step
    metrics_host = torch.empty(2, 8, dtype=torch.float32).pin_memory()
    read_back = [torch.cuda.Event() for _ in range(2)]
    last_read = [None]  # slot whose metrics the host has not read yet
    host_loss = [0.0]

    def enqueue_h2d(slot: int):
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed[slot])
            for k in host:
                if k == "hidden" and world > 1:
                    sl = slice(rank * rows, (rank + 1) * rows)
                    staging[slot][k][sl].copy_(host[k][sl], non_blocking=True)
                else:
                    staging[slot][k].copy_(host[k], non_blocking=True)
            if gather == "ce":
                hdl = sym_handles[slot]
                hdl.barrier(channel=0)  # every rank's rows of this slot are in place (and nobody still computes on it)
                dst = staging[slot]["hidden"]
                for i in range(1, world):  # staggered: at any moment every rank reads from a different peer
                    r = (rank + i) % world
                    sl = slice(r * rows, (r + 1) * rows)
                    dst[sl].copy_(peer_views[slot][r][sl], non_blocking=True)  # contiguous same-dtype: cudaMemcpyAsync
                hdl.barrier(channel=0)  # all pulls done: a rank may overwrite its rows at the slot's next use
            ready[slot].record(copy_stream)

    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    seg = {"fwd": [], "loss": [], "bwd": []}

    def step(e2e: bool, record: bool, slot: int = 0, prefetch: bool = False):
        if e2e:
            cur = torch.cuda.current_stream()
            cur.wait_event(ready[slot])
            if prefetch:
                enqueue_h2d(slot ^ 1)
            src = staging[slot]
            if gather == "nccl":
                dist.all_gather_into_tensor(src["hidden"], src["hidden"][rank * rows:(rank + 1) * rows], group=group)
        else:
            src = dict(hidden=hidden_d, labels=labels_d, old=old_d, ref=ref_d, adv=adv_d, mask=mask_d)
        hidden = src["hidden"].detach().requires_grad_(True)
        weight.grad = None
        marks = [ev() for _ in range(4)] if record else None
        if record:
            marks[0].record()
        lp, ent = linear_cross_entropy(hidden, weight, src["labels"], 1.0, "none", group)
        if record:
            marks[1].record()
        loss, metrics = core_algos.fused_actor_loss(src["old"], lp.view(1, -1), src["adv"], src["mask"],
                                                    entropy=ent.view(1, -1), ref_log_prob=src["ref"], **LOSS_CFG)
        if record:
            marks[2].record()
        loss.backward()  # world > 1: d_hidden comes back already summed over the ranks (kernels.set_reduce_d_hidden)
        if record:
            marks[3].record()
            seg["_marks"].append(marks)
        if e2e:
            consumed[slot].record()
            # the host reads EVERY step's loss / metrics, one step behind the launches: step i's 32 bytes are waited
            # for after step i+1 has been enqueued, so the GPU never idles while Python builds the next step
            metrics_host[slot].copy_(metrics, non_blocking=True)
            read_back[slot].record()
            if last_read[0] is not None:
                read_back[last_read[0]].synchronize()
                host_loss[0] = float(metrics_host[last_read[0]][0])
            last_read[0] = slot
        return metrics

    def timed(e2e: bool, steps: int, record: bool):
        seg["_marks"] = []
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = ev(), ev()
        e0.record()
        if e2e:
            enqueue_h2d(0)  # step 0's inputs: inside the timed region, not hidden behind anything
        for i in range(steps):
            step(e2e, record, slot=i & 1, prefetch=i + 1 < steps)
        if e2e and last_read[0] is not None:  # the last step's metrics are read inside the timed region too
            read_back[last_read[0]].synchronize()
            host_loss[0] = float(metrics_host[last_read[0]][0])
            last_read[0] = None
        e1.record()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    for _ in range(max(args.warmup, 3)):
        step(False, False)
    enqueue_h2d(0)
    step(True, False)
    torch.cuda.synchronize()
    last_read[0] = None

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = _lib.launch_count()
    ms_dev = timed(False, args.steps, True)
    launches = _lib.launch_count() - launches0
    marks = seg["_marks"]
    ms_e2e = timed(True, args.steps, False)
    clocks = sampler.stop() if rank == 0 else None

    fwd_ms = sum(m[0].elapsed_time(m[1]) for m in marks) / len(marks)
    loss_ms = sum(m[1].elapsed_time(m[2]) for m in marks) / len(marks)
    bwd_ms = sum(m[2].elapsed_time(m[3]) for m in marks) / len(marks)

    if rank == 0:
        peaks = _peaks()
        flops_fwd = 2.0 * T * H * v_local  # per GPU
        fwd_tf = flops_fwd / (fwd_ms * 1e-3) / 1e12
        step_useful_tf = 3 * flops_fwd / (ms_dev * 1e-3) / 1e12
        step_exec_tf = 4 * flops_fwd / (ms_dev * 1e-3) / 1e12
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            n_cpu = args.cpu_tokens or 4096
            tps, dt = cpu_reference_sample(n_cpu, threads)
            cpu = {"value": tps, "unit": "tokens/s", "cores": threads, "kind": "port",
                   "sample": f"{n_cpu} tokens at full H and V, fp32 math on bf16 inputs, {dt:.1f} s"}
        line = {
            "metric": METRIC, "value": T / (ms_dev * 1e-3), "unit": "tokens/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_dev, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": workload_config(T, world, exchange), "tp_reduce": reduce_how,
            "e2e": {"value": T / (ms_e2e * 1e-3), "unit": "tokens/s", "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": 32 * world,
                    "how": "pinned host -> device on a copy stream, double-buffered (step i+1's inputs copy under step "
                           "i's compute; every copy is inside the timed region)"
                           + ("" if world == 1 else f"; each rank copies T/{world} rows of hidden and the ranks "
                              + ("pull the other rows peer-to-peer with copy-engine DMA over NVLink (symmetric staging "
                                 "slots, no SM work)" if gather == "ce" else
                                 "all-gather the replicated tensor over NVLink (NCCL)"))
                           + "; metrics[8] of every step read by the host, one step behind the launches"},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {
                "kernel": "tc_gemm_kernel<MODE_FWD> (tcgen05 lm-head forward with fused online-softmax epilogue)",
                "bound": "tensor", "achieved": fwd_tf, "peak": peaks["sustained"], "unit": "TFLOP/s",
                "frac": fwd_tf / peaks["sustained"], "traffic": _traffic(T, H, v_local),
                "traffic_unit": "bytes of DRAM read+write per launch (ncu --set full, profiles/roofline_traffic.json); "
                                "algorithmic operand bytes = 2*(T*H + V*H) + 16*S*T",
                "peak_source": peaks["source"] + ", sustained bf16 figure (kernel timed inside a long step)",
                "frac_of_burst_peak": fwd_tf / peaks["burst"],
                "algorithmic_flops_per_launch": flops_fwd, "launch_ms": fwd_ms,
            },
            "roofline_step": {
                "useful_tflops": step_useful_tf, "executed_tflops": step_exec_tf,
                "useful_frac_of_sustained": step_useful_tf / peaks["sustained"],
                "useful_frac_of_burst": step_useful_tf / peaks["burst"],
                "executed_frac_of_sustained": step_exec_tf / peaks["sustained"],
                "segments_ms": {"lm_head_fwd": fwd_ms, "ppo_loss": loss_ms, "lm_head_bwd(+allreduce)": bwd_ms},
                "note": "useful = 6*T*H*V/N per GPU (fwd + dH + dW), executed = 8*T*H*V/N (tile recompute in the backward)",
            },
            "cpu_baseline": cpu,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--tokens", type=int, default=TOKENS)
    ap.add_argument("--cpu-tokens", type=int, default=0,
                    help="token sample of the CPU legs (0 = 4096 for cpu_baseline, 2048 per step for --impl reference)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--tp-reduce", choices=["auto", "nccl", "multimem", "overlap"], default="auto",
                    help="N > 1: d_hidden all-reduce after the GEMMs (nccl), under the d_weight GEMM (overlap), chosen by the "
                         "cost model (auto, default) or fused into the d_hidden GEMM (multimem.red)")
    ap.add_argument("--e2e-gather", choices=["ce", "nccl"], default="ce",
                    help="N>1 end-to-end leg: how the replicated hidden is assembled from the per-rank H2D shards")
    ap.add_argument("--tp-exchange", choices=["auto", "nccl", "nvlink"], default="auto",
                    help="N > 1: fused NVLink exchange (default when NVSwitch multicast is available) or NCCL")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
