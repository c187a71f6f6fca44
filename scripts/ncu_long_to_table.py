"""Pivot an `ncu --metrics ... --csv` log (one line per launch and metric) into one line per launch:
    python scripts/ncu_long_to_table.py in.csv > table.csv"""
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
h = rows[0]
ki, mi, ui, vi, ii = (h.index(k) for k in ("Kernel Name", "Metric Name", "Metric Unit", "Metric Value", "ID"))
launch, metrics, units = {}, [], {}
for r in rows[1:]:
    launch.setdefault(r[ii], {"Kernel Name": r[ki]})[r[mi]] = r[vi].replace(",", "")
    if r[mi] not in metrics:
        metrics.append(r[mi])
        units[r[mi]] = r[ui]
w = csv.writer(sys.stdout)
w.writerow(["Kernel Name"] + metrics)
w.writerow([""] + [units[m] for m in metrics])
for v in launch.values():
    w.writerow([v["Kernel Name"]] + [v.get(m, "") for m in metrics])
