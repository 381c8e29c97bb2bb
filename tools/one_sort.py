"""Run a few device sorts of one (kind, n): the command ncu wraps.  Dev tool."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import radix_sort_b200 as rs
from radix_sort_b200 import distr, verify

ap = argparse.ArgumentParser()
ap.add_argument("--kind", default="u32")
ap.add_argument("--n", type=float, default=2 ** 26)
ap.add_argument("--variant", type=int, default=-1)
ap.add_argument("--reps", type=int, default=2)
ap.add_argument("--distr", default="UNIFORM")
ap.add_argument("--param", type=float, default=0.0)
a = ap.parse_args()
k = rs.Kind[a.kind.upper()]
n = int(a.n)
dev = torch.device("cuda:0")
rs.set_variant(rs.key_bytes(k), a.variant)
src = distr.empty_keys(n, k, dev)
distr.generate(src, k, distr.Distr[a.distr], a.param)
fp = verify.fingerprint(src, k)
keys = torch.empty_like(src)
ws = torch.empty(rs.temp_bytes(n, k), dtype=torch.uint8, device=dev)
for _ in range(a.reps):
    keys.copy_(src)
    rs.radix_sort(keys, k, workspace=ws)
torch.cuda.synchronize()
assert verify.count_unsorted(keys, k) == 0 and verify.fingerprint(keys, k) == fp
print("ok", a.kind, n)
