"""Aggregate an `ncu --page source --csv` dump by SASS opcode: executed count, stall samples.  Dev tool."""
import csv
import re
import sys
from collections import defaultdict

path = sys.argv[1]
rows = list(csv.reader(open(path)))
hdr = rows[1]
ia, isrc, isamp, iexec = hdr.index("Address"), hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
ops = defaultdict(lambda: [0, 0])
tot_s = tot_e = 0
for r in rows[2:]:
    if len(r) <= iexec or not r[ia].startswith("0x"):
        continue
    m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[isrc])
    if not m:
        continue
    op = m.group(2)
    op = ".".join(op.split(".")[:2])
    e, s = int(r[iexec] or 0), int(r[isamp] or 0)
    ops[op][0] += e
    ops[op][1] += s
    tot_e += e
    tot_s += s
print(f"total executed {tot_e}  samples {tot_s}")
for op, (e, s) in sorted(ops.items(), key=lambda kv: -kv[1][1])[:int(sys.argv[2]) if len(sys.argv) > 2 else 25]:
    print(f"{op:22s} exec {e:>12d} ({100*e/tot_e:5.1f}%)  samples {s:>8d} ({100*s/max(1,tot_s):5.1f}%)")
