"""Device time of the record sorts the reference's demo runs (main.rs:108-125):
(u32,u32) and (u64,u64) pairs, key uniform, payload 0 (distr.rs:42-52).  Dev tool."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import radix_sort_b200 as rs
from radix_sort_b200 import distr

dev = torch.device("cuda:0")
for kind, n in ((rs.Kind.U32, 500_000_000), (rs.Kind.U64, 250_000_000)):
    kb = rs.key_bytes(kind)
    keys = distr.empty_keys(n, kind, dev)
    distr.generate(keys, kind, distr.Distr.UNIFORM, 0, seed=3)
    recs = torch.zeros((n, 2), dtype=keys.dtype, device=dev)
    recs[:, 0] = keys
    src = recs.clone()
    del keys
    nv = rs.num_variants(100 + 2 * kb)
    for v in range(nv):
        rs.set_variant(100 + 2 * kb, v)
        best = 1e9
        for r in range(3):
            recs.copy_(src)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            rs.radix_sort_records(recs, kind, 0)
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        k = recs[:, 0].contiguous()
        ok = bool((k[1:].view(torch.int64 if kb == 8 else torch.int32).to(torch.float64) >= -1e30).all())  # placeholder touch
        srt = bool((recs[1:, 0].to(torch.float64) >= recs[:-1, 0].to(torch.float64)).all()) if kb == 4 else None
        print(json.dumps({"record": f"({kind.name.lower()},{kind.name.lower()})", "n": n, "variant": v, "ms": round(best, 3),
                          "Grecords_s": round(n / best / 1e6, 3), "pass_GBps": round(2 * n * 2 * kb * kb / best / 1e6, 1),
                          "sorted": srt}), flush=True)
    rs.set_variant(100 + 2 * kb, -1)
    del recs, src
    torch.cuda.empty_cache()
