"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel totals,
shares, and the K3 share of one sort.  usage: launch_summary.py <launches.csv>"""
import csv
import re
import sys
from collections import defaultdict

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if not l.startswith("==")) if r]
hdr = rows[0]
ik, iv, ig = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size")
iu = hdr.index("Metric Unit")
agg = defaultdict(lambda: [0.0, 0])
by_grid = defaultdict(lambda: [0.0, 0])
n = 0
for r in rows[1:]:
    v = float(r[iv].replace(",", ""))
    us = v / 1000.0 if r[iu] in ("ns", "nsecond") else v
    name = re.sub(r"\(.*", "", r[ik]).replace("b200::", "")
    agg[name][0] += us
    agg[name][1] += 1
    by_grid[(name, r[ig])][0] += us
    by_grid[(name, r[ig])][1] += 1
    n += 1
tot = sum(a[0] for a in agg.values())
print(f"# launches captured: {n}; total device time in them: {tot / 1000:.3f} ms")
for name, (us, c) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    print(f"{100 * us / tot:6.2f}%  {us / 1000:10.3f} ms {c:5d} launches  avg {us / c:10.1f} us  {name[:110]}")
print("# by grid size (top 8):")
for (name, g), (us, c) in sorted(by_grid.items(), key=lambda kv: -kv[1][0])[:8]:
    print(f"   grid {g:>22s} {c:4d} launches  avg {us / c:10.1f} us  {name[:90]}")
