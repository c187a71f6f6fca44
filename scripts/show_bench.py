"""Print the essentials of bench.py JSON lines: python scripts/show_bench.py file..."""
import json, sys
for f in sys.argv[1:]:
    try:
        for line in open(f):
            if not line.startswith("{"):
                continue
            d = json.loads(line)
            if "ms_per_step" not in d:
                print(f, json.dumps(d)[:400]); continue
            rk = {k: round(v["ms_per_step"], 2) for k, v in (d.get("roofline_kernels") or {}).items()}
            seg = {k: (round(v, 2) if v is not None else None) for k, v in ((d.get("roofline_step") or {}).get("segments_ms") or {}).items()}
            mem = d.get("memory") or {}
            print(f"{f}: N={d['n_gpus']} {d['ms_per_step']:.2f} ms/step {d['value']:.0f} tok/s e2e {(d.get('e2e') or {}).get('ms_per_step') or float('nan'):.2f} ms "
                  f"reduce={d.get('tp_reduce')} parity={(d.get('parity') or {}).get('ok')} chunks={mem.get('bwd_chunks')} "
                  f"({mem.get('bwd_chunking')})\n   kernels {rk}\n   segments {seg}\n   clocks {d.get('clocks')}\n   roofline frac {(d.get('roofline') or {}).get('frac')} "
                  f"useful/burst {(d.get('roofline_step') or {}).get('useful_frac_of_burst')}")
    except Exception as e:  # noqa: BLE001
        print(f, "ERR", e)
