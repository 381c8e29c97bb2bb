"""Aggregate an `ncu --page source --csv` dump by CUDA source line (needs -lineinfo): executed warp-instructions,
shared wavefronts, stall samples.  usage: ncu_lines.py <csv> <kernel-substring> [top]"""
import csv
import re
import sys
from collections import defaultdict

rows = list(csv.reader(open(sys.argv[1])))
want = sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
# the dump holds several kernels, each introduced by a "Kernel Name" row then a header row
i = 0
agg = defaultdict(lambda: [0, 0, 0, []])
tot = [0, 0, 0]
while i < len(rows):
    if rows[i] and rows[i][0] == "Kernel Name":
        name = rows[i][1]
        hdr = rows[i + 1]
        i += 2
        use = want in name
        if use:
            print("kernel:", name[:120])
        ia, isrc, isamp, iexec = hdr.index("Address"), hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
        iwf = hdr.index("L1 Wavefronts Shared") if "L1 Wavefronts Shared" in hdr else None
        while i < len(rows) and not (rows[i] and rows[i][0] == "Kernel Name"):
            r = rows[i]
            i += 1
            if not use or len(r) <= iexec or not r[ia].startswith("0x"):
                continue
            m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[isrc])
            op = ".".join(m.group(2).split(".")[:2]) if m else "?"
            e, s = int(r[iexec] or 0), int(r[isamp] or 0)
            w = int(r[iwf] or 0) if iwf is not None else 0
            a = agg[op]
            a[0] += e; a[1] += s; a[2] += w
            tot[0] += e; tot[1] += s; tot[2] += w
        if use:
            break
    else:
        i += 1
print(f"total: executed {tot[0]}  samples {tot[1]}  shared wavefronts {tot[2]}")
for op, (e, s, w, _) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{op:22s} exec {e:>11d} ({100*e/max(1,tot[0]):5.1f}%)  samples {s:>7d} ({100*s/max(1,tot[1]):5.1f}%)  smem wavefronts {w:>10d}")
