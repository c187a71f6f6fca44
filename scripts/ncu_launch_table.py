"""Per-kernel table from an ncu --csv launch list with gpu__time_duration.sum + dram bytes:
    python scripts/ncu_launch_table.py launches.csv [min_MB]   -> kernel, grid, time, DRAM MB, GB/s (largest launch of each)"""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
min_mb = float(sys.argv[2]) if len(sys.argv) > 2 else 50.0
hdr = rows[0]
ki, mi, vi, ii, gi = (hdr.index(k) for k in ("Kernel Name", "Metric Name", "Metric Value", "ID", "Grid Size"))
d = collections.OrderedDict()
for r in rows[1:]:
    d.setdefault(r[ii], {"k": r[ki].replace("void vpg::", "")[:64], "g": r[gi]})[r[mi]] = float(r[vi].replace(",", ""))
seen = collections.OrderedDict()
for v in d.values():
    seen.setdefault((v["k"], v["g"]), []).append(v)
for (k, g), vs in seen.items():
    v = max(vs, key=lambda x: x.get("dram__bytes_read.sum", 0))
    t = v["gpu__time_duration.sum"] / 1e3
    b = v.get("dram__bytes_read.sum", 0) + v.get("dram__bytes_write.sum", 0)
    if b > min_mb * 1e6:
        print(k.ljust(66), g.ljust(14), "%8.1f us %8.1f MB %6.0f GB/s" % (t, b / 1e6, b / t / 1e3))
