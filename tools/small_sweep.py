"""Device time of small sorts: the one-CTA small sort against the K1-K2-K3 sequence.  Dev tool."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import radix_sort_b200 as rs
from radix_sort_b200 import distr

dev = torch.device("cuda:0")
for kname in ("u8", "u32", "u64", "u128"):
    k = rs.Kind[kname.upper()]
    for n in (100, 1000, 4096):
        src = distr.empty_keys(n, k, dev)
        distr.generate(src, k, distr.Distr.UNIFORM, 0, seed=1)
        keys = torch.empty_like(src)
        ws = torch.empty(rs.temp_bytes(n, k), dtype=torch.uint8, device=dev)
        row = {"kind": kname, "n": n}
        for small in (True, False):
            rs.debug_small_sort(small)
            best = 1e9
            for r in range(20):
                keys.copy_(src)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                rs.radix_sort(keys, k, workspace=ws)
                e1.record()
                torch.cuda.synchronize()
                best = min(best, e0.elapsed_time(e1))
            row["one_cta_us" if small else "k1_k2_k3_us"] = round(best * 1e3, 1)
        rs.debug_small_sort(True)
        print(json.dumps(row), flush=True)
