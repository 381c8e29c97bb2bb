"""Wall time of the host drop-in path (numpy array in, sorted in place) by slice size.  Dev tool."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import radix_sort_b200 as rs

rng = np.random.default_rng(1)
for dt in (np.uint32, np.uint64):
    for n in (100, 1000, 4096, 10_000, 100_000, 1_000_000, 10_000_000):
        src = rng.integers(0, 2**32, size=n, dtype=np.uint64).astype(dt)
        a = src.copy()
        rs.radix_sort(a)  # warm-up: pools, first launch
        best = 1e9
        for _ in range(10 if n <= 1_000_000 else 3):
            a[:] = src
            t0 = time.perf_counter()
            rs.radix_sort(a)
            best = min(best, time.perf_counter() - t0)
        t0 = time.perf_counter()
        b = np.sort(src, kind="stable")
        cpu = time.perf_counter() - t0
        assert np.array_equal(a, b)
        print(json.dumps({"dtype": np.dtype(dt).name, "n": n, "host_path_us": round(best * 1e6, 1),
                          "Mkeys_s": round(n / best / 1e6, 1), "numpy_stable_sort_us": round(cpu * 1e6, 1)}), flush=True)
