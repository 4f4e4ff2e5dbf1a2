"""Sums an ncu launch list (--metrics gpu__time_duration.sum --csv) by kernel name: launches, total us, share.
Usage: python tools/launch_summary.py launches.csv [skip_first_n]"""
import csv
import re
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = next(r for r in rows if "Kernel Name" in r)
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
data = [r for r in rows if r is not hdr and len(r) == len(hdr) and r[ki] != "Kernel Name"]
half = len(data) // 2 if len(sys.argv) < 3 else int(sys.argv[2])
data = data[half:]                                     # the second build
agg = {}
for r in data:
    name = re.sub(r"<.*", "", r[ki].split("(")[0]).split("::")[-1]
    v = float(r[vi].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r[ui], 1.0)
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
print(f"# {sys.argv[1]}: {len(data)} launches, {tot:.0f} us of kernels")
for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"  {k:45s} n={a[0]:4d}  {a[1]:9.1f} us  {100 * a[1] / tot:5.1f} %")
