"""Split an `ncu --page source --csv` dump of one K3 kernel into the phases between
its barriers (SASS address order, executed instructions only): stall samples,
executed instructions and shared wavefronts per phase, and the heaviest instructions.
usage: ncu_phases.py <csv> <kernel-substring> [instance] [top]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
want = sys.argv[2]
inst = int(sys.argv[3]) if len(sys.argv) > 3 else 0
top = int(sys.argv[4]) if len(sys.argv) > 4 else 5
idx = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name" and want in r[1]]
i = idx[inst]
name = rows[i][1]
hdr = rows[i + 1]
i += 2
ia, isrc, isamp, iexec = hdr.index("Address"), hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
iwf = hdr.index("L1 Wavefronts Shared")
print("kernel:", name[:140])
ins = []
while i < len(rows) and not (rows[i] and rows[i][0] == "Kernel Name"):
    r = rows[i]
    i += 1
    if len(r) <= iexec or not r[ia].startswith("0x"):
        continue
    e, s, w = int(r[iexec] or 0), int(r[isamp] or 0), int(r[iwf] or 0)
    if e == 0 and s == 0:
        continue
    ins.append((s, e, w, r[ia][-5:], r[isrc].strip()))
tot = sum(x[0] for x in ins)
tw = sum(x[2] for x in ins)
print(f"total samples {tot}, shared wavefronts {tw}, executed {sum(x[1] for x in ins)}")
phase = []
k = 0
for x in ins:
    phase.append(x)
    if "BAR." in x[4] or "EXIT" in x[4]:
        s = sum(y[0] for y in phase)
        if s * 300 >= tot:
            print(f"--- phase {k}: samples {s} ({100*s/tot:.1f}%)  exec {sum(y[1] for y in phase)}  smem wf {sum(y[2] for y in phase)} ({100*sum(y[2] for y in phase)/max(1,tw):.1f}%)  [{phase[0][3]}..{phase[-1][3]}] first: {phase[0][0]} samples")
            for y in sorted(phase, reverse=True)[:top]:
                print(f"      {y[0]:>7d} {100*y[0]/tot:5.1f}%  exec {y[1]:>9d} wf {y[2]:>9d}  {y[3]} {y[4][:80]}")
        phase = []
        k += 1
