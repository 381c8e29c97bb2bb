"""Whole-sort device time of 1- and 2-byte key sorts, counting path vs onesweep passes.  Dev tool."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import radix_sort_b200 as rs
from radix_sort_b200 import distr, verify

dev = torch.device("cuda:0")
for kname in ("u8", "i8", "u16", "i16"):
    k = rs.Kind[kname.upper()]
    for n in (1 << 24, 1 << 27, 1 << 30, (1 << 30) + 12345):
        for dname, param in (("UNIFORM", 0.0), ("GEOMETRIC", 0.05)):
            src = distr.empty_keys(n, k, dev)
            distr.generate(src, k, distr.Distr[dname], param, seed=0xB2000006)
            fp = verify.fingerprint(src, k)
            keys = torch.empty_like(src)
            ws = torch.empty(rs.temp_bytes(n, k), dtype=torch.uint8, device=dev)
            for fill in (True, False):
                rs.debug_narrow_fill(fill)
                best = None
                for r in range(4):
                    keys.copy_(src)
                    torch.cuda.synchronize()
                    a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a0.record()
                    rs.radix_sort(keys, k, workspace=ws)
                    a1.record()
                    torch.cuda.synchronize()
                    ms = a0.elapsed_time(a1)
                    if r and (best is None or ms < best):
                        best = ms
                ok = verify.count_unsorted(keys, k) == 0 and verify.fingerprint(keys, k) == fp
                kb = rs.key_bytes(k)
                print(json.dumps({"kind": kname, "n": n, "distr": dname, "path": "histogram+fill" if fill else "onesweep passes", "ok": ok,
                                  "ms": round(best, 4), "gkeys_s": round(n / best / 1e6, 1),
                                  "GBps_2s_per_key": round(2 * kb * n / best / 1e6, 1)}), flush=True)
            rs.debug_narrow_fill(True)
            del src, keys, ws
            torch.cuda.empty_cache()
