"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv) by kernel: python scripts/launch_summary.py in.csv "<command>" > out.txt"""
import csv, sys
from collections import defaultdict
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
h = rows[0]
kn, mv, mu = h.index('Kernel Name'), h.index('Metric Value'), h.index('Metric Unit')
tot, cnt = defaultdict(float), defaultdict(int)
for r in rows[1:]:
    v = float(r[mv].replace(',', ''))
    v *= {'ns': 1e-6, 'us': 1e-3, 'usecond': 1e-3, 'ms': 1.0, 'msecond': 1.0, 'nsecond': 1e-6}.get(r[mu], 1e-6)
    tot[r[kn]] += v
    cnt[r[kn]] += 1
s = sum(tot.values())
print(sys.argv[2] if len(sys.argv) > 2 else '')
print('per-launch times are cold-cache and serialised: compare SHARES')
print(f"{'kernel':92s} {'launches':>8s} {'total_ms':>10s} {'share':>7s}")
for k in sorted(tot, key=tot.get, reverse=True):
    print(f"{k[:90]:92s} {cnt[k]:8d} {tot[k]:10.2f} {100*tot[k]/s:6.1f}%")
