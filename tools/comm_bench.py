"""Single-process multi-GPU sort (b200sort_mgpu_keys_device) on all visible GPUs:
host wall clock around the blocking call, shards resident, buffers warm.  Dev tool;
the contract's multi-GPU numbers come from bench.py (one process per GPU)."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import radix_sort_b200 as rs
from radix_sort_b200 import distr, mgpu, verify

n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 1_000_000_000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
K = rs.Kind.U64
g = torch.cuda.device_count()
shards = []
for r in range(g):
    t = distr.empty_keys(n, K, torch.device(f"cuda:{r}"))
    with torch.cuda.device(r):
        distr.generate(t, K, distr.Distr.UNIFORM, 0, seed=0xB2000002, index_base=r * n)
    shards.append(t)
torch.cuda.synchronize()
comm = mgpu.Communicator(list(range(g)))
best = 1e9
for i in range(reps + 1):
    t0 = time.perf_counter()
    outs = comm.sort(shards, K, copy=False)
    dt = time.perf_counter() - t0
    if i:
        best = min(best, dt)
ok = True
for r, o in enumerate(outs):
    with torch.cuda.device(r):
        ok = ok and verify.count_unsorted(o, K) == 0
bounds_ok = all(int(outs[r][-1:].cpu().numpy()[0]) <= int(outs[r + 1][:1].cpu().numpy()[0]) for r in range(g - 1) if len(outs[r]) and len(outs[r + 1]))
print(json.dumps({"what": "b200sort_mgpu_keys_device, one process", "n_gpus": g, "keys_per_gpu": n, "kind": "u64",
                  "best_ms": round(best * 1e3, 3), "gkeys_per_s": round(g * n / best / 1e9, 2),
                  "n_out": [len(o) for o in outs], "sorted": bool(ok), "ranges_ordered": bool(bounds_ok)}))
comm.close()
