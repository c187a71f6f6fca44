"""Timeline of one vocab-parallel step from the per-rank kernel traces `bench.py --trace PREFIX` wrote (torch.profiler /
CUPTI; timestamps share the host clock, so rank skew is visible).  Usage: python scripts/trace_summary.py PREFIX [step]
Prints, for the chosen step (default: the second of the three traced), every kernel longer than 20 us on every rank:
start and end relative to the earliest forward-kernel start of that step, and a per-rank line of idle gaps."""
import glob, json, sys


def short(n):
    for key, tag in (("tc_gemm_kernel<0,", "lm_head_fwd"), ("tc_gemm_kernel<1,", "dlogits_recompute"),
                     ("tc_gemm_kernel<2, 1, 1", "d_weight_gemm"), ("tc_gemm_kernel<2, 0, 1", "d_hidden_gemm"),
                     ("lce_finalize", "finalize"), ("ppo_loss", "ppo_loss")):
        if key in n:
            return tag
    return n.split("(")[0].replace("void ", "").replace("vpg::", "")[:56]


def main():
    prefix = sys.argv[1]
    which = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    ranks = []
    for f in sorted(glob.glob(prefix + ".rank*.json")):
        ranks.append(json.load(open(f)))
    ranks.sort(key=lambda r: r["rank"])
    rows, t0 = [], None
    for r in ranks:
        ev = r["events"]
        # a step starts at a forward kernel: the k-th kernel whose name contains the forward's mode tag
        fwd_idx = [i for i, e in enumerate(ev) if "tc_gemm_kernel<0," in e["name"]]
        if len(fwd_idx) <= which:
            fwd_idx = [0]
            lo, hi = 0, len(ev)
        else:
            lo = fwd_idx[which]
            hi = fwd_idx[which + 1] if which + 1 < len(fwd_idx) else len(ev)
        rows.append((r["rank"], ev[lo:hi]))
        t0 = ev[lo]["ts_us"] if t0 is None else min(t0, ev[lo]["ts_us"])
    print(f"config {ranks[0]['config']} world {ranks[0]['world']} tp_reduce {ranks[0]['tp_reduce']}; times in ms from each "
          f"rank's own forward start (CUPTI clocks of different processes are not aligned to better than the profiler's "
          f"per-process start, so cross-rank offsets are not shown)")
    for rank, ev in rows:
        t0 = ev[0]["ts_us"]
        print(f"-- rank {rank}")
        prev_end = None
        busy = 0.0
        for e in ev:
            s, d = (e["ts_us"] - t0) / 1e3, e["dur_us"] / 1e3
            gap = "" if prev_end is None else f" gap {s - prev_end:+.3f}"
            if d > 0.02 or (prev_end is not None and s - prev_end > 0.05):
                print(f"   {s:8.3f} -> {s + d:8.3f}  ({d:7.3f})  {short(e['name'])}{gap}")
            prev_end = max(prev_end or 0, s + d)
            busy += d
        print(f"   step span {prev_end - (ev[0]['ts_us'] - t0) / 1e3:.3f} ms, kernel-busy {busy:.3f} ms")


if __name__ == "__main__":
    main()
