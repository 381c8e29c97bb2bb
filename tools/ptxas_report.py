"""Summarise `nvcc -Xptxas -v` output: registers / spills / smem per kernel.  Dev tool."""
import re
import subprocess
import sys

text = open(sys.argv[1]).read()
filt = sys.argv[2] if len(sys.argv) > 2 else ""
names = re.findall(r"Compiling entry function '(\S+)' for 'sm_100a'", text)
blocks = re.split(r"Compiling entry function '\S+' for 'sm_100a'", text)[1:]
dem = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
for name, blk in zip(dem, blocks):
    sp = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", blk)
    rg = re.search(r"Used (\d+) registers", blk)
    short = re.sub(r"b200::", "", name)
    short = re.sub(r"\(.*", "", short)
    if filt and filt not in short:
        continue
    print(f"regs {rg.group(1) if rg else '?':>3}  stack {sp.group(1):>4} spill {sp.group(2):>4}/{sp.group(3):<4} {short[:110]}")
