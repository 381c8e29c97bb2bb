"""Where does the end-to-end time of b200sort_keys_host go?  Raw pinned H2D / D2H
rates next to the plain and the pipelined host path.  Dev tool."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import radix_sort_b200 as rs
from radix_sort_b200 import _ffi, distr

n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 1_000_000_000
dev = torch.device("cuda:0")
K = rs.Kind.U64
src = distr.empty_keys(n, K, dev)
distr.generate(src, K, distr.Distr.UNIFORM, 0, seed=1)
host = torch.empty(n, dtype=torch.uint64, pin_memory=True)
d = torch.empty_like(src)


def timed(fn, reps=3):
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize()
        t = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t)
    return best * 1e3


host.copy_(src)
print("H2D 8n bytes  ms", round(timed(lambda: d.copy_(host, non_blocking=True)), 2))
print("D2H 8n bytes  ms", round(timed(lambda: host.copy_(d, non_blocking=True)), 2))
L = _ffi.lib()
for name, thr in (("plain", 2**63), ("pipelined", 0)):
    rs.set_host_pipeline_min_bytes(thr)
    best = 1e9
    for _ in range(3):
        host.copy_(src)
        torch.cuda.synchronize()
        t = time.perf_counter()
        _ffi.check(L.b200sort_keys_host(host.data_ptr(), n, int(K)))
        best = min(best, time.perf_counter() - t)
    print(name, "keys_host ms", round(best * 1e3, 2), "Gkeys/s", round(n / best / 1e9, 3))

# pageable host memory (what a Rust Vec is): the same call on an unpinned numpy array
import numpy as np
rs.set_host_pipeline_min_bytes(64 << 20)
m = min(n, 500_000_000)
a = src[:m].cpu().numpy().copy()
for _ in range(2):
    b = a.copy()
    t = time.perf_counter()
    _ffi.check(L.b200sort_keys_host(b.ctypes.data, m, int(K)))
    dt = time.perf_counter() - t
print("pageable keys_host ms", round(dt * 1e3, 2), "Gkeys/s", round(m / dt / 1e9, 3), "(n =", m, ")")

# pageable (key, value) records: the demo's (u32,u32) and (u64,u64) shapes (main.rs:108-125)
for kind, rec_dtype, mrec in ((rs.Kind.U32, np.uint32, min(n // 2, 500_000_000)), (rs.Kind.U64, np.uint64, min(n // 4, 250_000_000))):
    rng = np.random.default_rng(3)
    recs = rng.integers(0, np.iinfo(rec_dtype).max, size=(mrec, 2), dtype=rec_dtype)
    for _ in range(2):
        b = recs.copy()
        t = time.perf_counter()
        rs.radix_sort_records(b, kind, key_off=0)
        dt = time.perf_counter() - t
    ok = bool((b[1:, 0] >= b[:-1, 0]).all())
    print("pageable records", rec_dtype.__name__, "ms", round(dt * 1e3, 2), "Grec/s", round(mrec / dt / 1e9, 3), "sorted", ok, "(n =", mrec, ")")
