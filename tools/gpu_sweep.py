"""On-GPU sweep of kernel variants: per-phase CUDA-event timings of the device
sort for each (kind, n, variant).  Development tool; prints one JSON per line."""
import argparse
import json
import sys
import os

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch

import radix_sort_b200 as rs
from radix_sort_b200 import distr, verify



def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--kinds", default="u32,u64")
    ap.add_argument("--n", default="268435456")
    ap.add_argument("--variants", default="all")
    ap.add_argument("--distr", default="UNIFORM")
    ap.add_argument("--param", type=float, default=0.0)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--flags", type=int, default=0)  # b200sort flags (FLAG_NO_SKIP)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    peak = 6552.0
    for kname in a.kinds.split(","):
        k = rs.Kind[kname.upper()]
        kb = rs.key_bytes(k)
        for n in [int(float(x)) for x in a.n.split(",")]:
            src = distr.empty_keys(n, k, dev)
            distr.generate(src, k, distr.Distr[a.distr], a.param, seed=0xB2000001)
            fp = verify.fingerprint(src, k) if a.check else None
            keys = torch.empty_like(src)
            ws = torch.empty(rs.temp_bytes(n, k), dtype=torch.uint8, device=dev)
            variants = range(rs.num_variants(kb)) if a.variants == "all" else [int(v) for v in a.variants.split(",")]
            for v in variants:
                rs.set_variant(kb, v)
                best = None
                for r in range(a.reps + 1):
                    keys.copy_(src)
                    torch.cuda.synchronize()
                    t = rs.sort_timed(keys, k, workspace=ws, flags=a.flags)
                    if r == 0:
                        continue  # warm-up
                    if best is None or t["total_ms"] < best["total_ms"]:
                        best = t
                ok = None
                if a.check:
                    ok = verify.count_unsorted(keys, k) == 0 and verify.fingerprint(keys, k) == fp
                passes = [p for p in best["pass_ms"] if p > 0.002]
                pass_ms = sum(passes) / max(1, len(passes))
                out = {
                    "kind": kname, "n": n, "variant": v, "distr": a.distr, "flags": a.flags, "ok": ok,
                    "total_ms": round(best["total_ms"], 4), "hist_ms": round(best["hist_ms"], 4),
                    "scan_ms": round(best["scan_ms"], 4), "copy_back_ms": round(best["copy_back_ms"], 4),
                    "pass_ms": [round(p, 4) for p in best["pass_ms"]],
                    "gkeys_s": round(n / best["total_ms"] / 1e6, 3),
                    "pass_gbs": round(2 * n * kb / pass_ms / 1e6, 1) if passes else None,
                    "pass_frac_of_measured": round(2 * n * kb / pass_ms / 1e6 / peak, 3) if passes else None,
                    "hist_gbs": round(n * kb / best["hist_ms"] / 1e6, 1),
                }
                print(json.dumps(out), flush=True)
            rs.set_variant(kb, -1)
            del src, keys, ws
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
