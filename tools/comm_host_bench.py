"""b200sort_mgpu_keys_host on all visible GPUs: one pinned (and one pageable) host
slice of n u64 keys, host wall clock around the blocking call.  Dev tool."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import radix_sort_b200 as rs
from radix_sort_b200 import distr, mgpu

n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 1_000_000_000
K = rs.Kind.U64
g = torch.cuda.device_count()
src = distr.empty_keys(n, K, torch.device("cuda:0"))
distr.generate(src, K, distr.Distr.UNIFORM, 0, seed=0xB2000002)
pinned = torch.empty(n, dtype=torch.uint64, pin_memory=True)
comm = mgpu.Communicator(list(range(g)))
out = {"what": "b200sort_mgpu_keys_host, one process", "n_gpus": g, "n": n, "kind": "u64"}
legs = [("pinned", pinned.numpy())]
if "pageable" in sys.argv[2:]:
    legs.append(("pageable", np.empty(n, dtype=np.uint64)))
for name, buf in legs:
    best = 1e9
    for i in range(3):
        buf[...] = src.cpu().numpy() if name == "pageable" else 0
        if name == "pinned":
            pinned.copy_(src)
            torch.cuda.synchronize()
        t0 = time.perf_counter()
        comm.sort_host(buf, K)
        dt = time.perf_counter() - t0
        if i:
            best = min(best, dt)
    out[name + "_ms"] = round(best * 1e3, 2)
    out[name + "_gkeys_per_s"] = round(n / best / 1e9, 3)
    out[name + "_sorted"] = bool(np.all(buf[1:] >= buf[:-1]))
print(json.dumps(out))
comm.close()
