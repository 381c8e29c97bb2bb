#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 radix sort (BASELINE.json metric:
sorted Gkeys/s; achieved HBM GB/s vs peak).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

A step = one full sort of the workload.  N = 1: BASELINE.json configs[1],
1 000 000 000 uniform random u64 keys resident in HBM (b200sort_keys_device).
N > 1 (launched under torchrun, one rank per GPU): the same 1e9 u64 keys PER
GPU, sharded (weak scaling; the shape of configs[4]): sample -> splitters ->
K7 partition fused with the exchange (peer stores over NVLink; NCCL all-to-all as
fallback) -> local LSD sort (radix_sort_b200/mgpu.py).
Every step starts from the same unsorted input (restored by a device copy
outside the timed region); the timed region of a step is bracketed by CUDA
events on the launching stream; rank time = sum over steps, job time = max
over ranks.  Inputs (8 GB per GPU) are far larger than the 126 MB L2, so no
flush is needed between steps.

e2e = the same metric through the drop-in C-ABI call b200sort_keys_host on a
pinned HOST buffer: H2D copy, sort, D2H copy all inside the timed region.

--impl reference times the reference's CPU algorithm (the C restatement under
oracle/, since no Rust toolchain exists on these boxes) on the host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_KEYS = 1_000_000_000          # BASELINE.json configs[1]
KIND_NAME = "u64"
KEY_BYTES = 8
DIGITS = 8
SEED = 0xB2000002
CPU_SAMPLE_KEYS = 100_000_000   # bounded CPU sample of the same workload (0.8 GB + 0.8 GB temp)


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:  # noqa: BLE001
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:  # noqa: BLE001
                self.proc.kill()
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx = max(mx, float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_baseline(n: int, steps: int = 1):
    """Time the oracle's restatement of the reference's production sort
    (mod.rs:61-176, native threads + 96-element write buffers) and of its rayon
    variant radix_sort4 (mod.rs:405-466) on all host cores."""
    import numpy as np
    from oracle import cpu_oracle
    from radix_sort_b200.distr import uniform_host
    from radix_sort_b200.api import Kind
    cores = cpu_oracle.available_parallelism()
    src = uniform_host(n, Kind.U64, seed=SEED)
    out = {}
    for name, fn in (("radix_sort", cpu_oracle.sort_mt), ("radix_sort4", cpu_oracle.sort4_mt)):
        best = None
        for _ in range(steps):
            a = src.copy()
            t0 = time.perf_counter()
            fn(a, "u64", cores)
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        assert bool(np.all(a[:-1] <= a[1:]))
        out[name] = n / best / 1e9
    return cores, out


def run_reference(args):
    """--impl reference: the reference's CPU path on this box's host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import numpy as np
    from oracle import cpu_oracle
    from radix_sort_b200.distr import uniform_host
    from radix_sort_b200.api import Kind
    cores = cpu_oracle.available_parallelism()
    n = CPU_SAMPLE_KEYS
    src = uniform_host(n, Kind.U64, seed=SEED)
    a = src.copy()
    for _ in range(args.warmup):
        a[...] = src
        cpu_oracle.sort_mt(a, "u64", cores)
    total = 0.0
    for _ in range(args.steps):
        a[...] = src
        t0 = time.perf_counter()
        cpu_oracle.sort_mt(a, "u64", cores)
        total += time.perf_counter() - t0
    assert bool(np.all(a[:-1] <= a[1:]))
    ms = total / args.steps * 1e3
    value = n / (ms * 1e-3) / 1e9
    sample = (f"{n} of the workload's {N_KEYS} uniform u64 keys per step; oracle_sort_mt = C restatement of "
              f"RadixSort::radix_sort (mod.rs:61-176), {cores} threads; the Rust crate itself cannot be built here")
    line = {
        "impl": "reference", "metric": "sorted Gkeys/s (u64)", "value": round(value, 4), "unit": "Gkeys/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": "uniform random u64 keys, 1B elements (BASELINE.json configs[1]); CPU arm runs a bounded sample",
                   "keys_per_step": n, "seed": SEED},
        "cpu_baseline": {"value": round(value, 4), "unit": "Gkeys/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": round(value, 4), "unit": "Gkeys/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def run_gpu(args):
    import numpy as np
    import torch
    import radix_sort_b200 as rs
    from radix_sort_b200 import distr, verify, mgpu, _ffi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the sort has no CPU fallback")
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29531")
        # stdout carries the ONE JSON line: NCCL's version banner (NCCL_DEBUG=VERSION on some boxes) goes to stderr
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)

    K = rs.Kind.U64
    n = args.n
    pristine = distr.empty_keys(n, K, dev)
    distr.generate(pristine, K, distr.Distr.UNIFORM, 0, seed=SEED, index_base=rank * n)
    keys = torch.empty_like(pristine)
    ws = torch.empty(rs.temp_bytes(int(n * 1.25) + 1024, K), dtype=torch.uint8, device=dev)
    fp_in = verify.fingerprint(pristine, K)
    # N > 1: fused K7 + exchange over peer memory (torch symmetric memory) unless
    # it cannot be set up here; then the NCCL all-to-all path, and the line says so
    exchange, exchange_note, sendbuf = None, "single GPU", None
    if world > 1:
        exchange_note = "nccl all_to_all_single"
        if args.exchange in ("auto", "fused"):
            try:
                exchange = mgpu.PeerExchange(int(n * 1.25) + 1024, KEY_BYTES, dev)
                exchange_note = "fused: K7 stores into peer receive buffers over NVLink (symmetric memory)"
            except Exception as e:  # noqa: BLE001
                if args.exchange == "fused":
                    raise
                exchange_note = f"nccl all_to_all_single (peer memory unavailable: {type(e).__name__}: {str(e)[:120]})"
        if exchange is None:
            sendbuf = torch.empty_like(pristine)
    # overlapped exchange: ranges cut in two, the second halves cross NVLink while the first are sorted
    # (measured slower than the one-piece exchange at N = 2 and N = 8 -- profiles/r01_bench_n*_overlap_*.json -- so opt-in)
    overlap = exchange is not None and args.overlap == "on" and 2 * world <= mgpu.MAX_RANKS
    ws2 = torch.empty(rs.temp_bytes(int(n * 0.75) + 1024, K), dtype=torch.uint8, device=dev) if overlap else None
    if overlap:
        exchange_note += "; exchange in two pieces, the second overlapped with the local sort of the first"

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def one_step():
        """returns (elapsed ms on the device, result tensor)"""
        keys.copy_(pristine)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        if world == 1:
            rs.radix_sort(keys, K, workspace=ws)
            out = keys
        else:
            out = mgpu.sort_sharded(keys, K, workspace=ws, send_buffer=sendbuf, exchange=exchange, overlap=overlap,
                                    workspace2=ws2).keys
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1), out

    for _ in range(max(args.warmup, 0)):
        one_step()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = rs.launch_count()
    barrier()
    total_ms = 0.0
    out = keys
    for _ in range(args.steps):
        ms, out = one_step()
        total_ms += ms
    barrier()
    launches = rs.launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else None
    phases = {}
    if world > 1:  # one extra, untimed step with per-phase device timings
        keys.copy_(pristine)
        barrier()
        out = mgpu.sort_sharded(keys, K, workspace=ws, send_buffer=sendbuf, exchange=exchange, timings=phases,
                                overlap=overlap, workspace2=ws2).keys
        barrier()

    # correctness of the last timed step: sorted + same multiset <=> bit-identical to the reference's output
    unsorted = verify.count_unsorted(out, K)
    fp_out = verify.fingerprint(out, K)
    ok = unsorted == 0
    if world == 1:
        ok = ok and fp_out == fp_in
    else:
        t = torch.tensor([fp_in[0], fp_in[1] % 2**63, fp_in[3] % 2**63, fp_out[0], fp_out[1] % 2**63, fp_out[3] % 2**63],
                         dtype=torch.int64, device=dev)
        # sums mod 2^63 of the per-rank fingerprints must agree before / after (xor term checked separately)
        both = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(both, t)
        tot = [sum(int(b[i]) for b in both) for i in range(6)]
        ok = ok and tot[0] == tot[3] and (tot[1] - tot[4]) % 2**63 == 0 and (tot[2] - tot[5]) % 2**63 == 0
        # shard boundaries: my first key >= previous rank's last key
        edge = torch.zeros(2, dtype=torch.int64, device=dev)
        if out.shape[0]:
            o64 = out.view(torch.int64)
            edge[0] = o64[0] ^ (-2**63)
            edge[1] = o64[-1] ^ (-2**63)
        edges = [torch.zeros_like(edge) for _ in range(world)]
        dist.all_gather(edges, edge)
        last = None
        for e, b in zip(edges, both):
            if int(b[3]) == 0:
                continue
            if last is not None and int(e[0]) < last:
                ok = False
            last = int(e[1])
    if not ok:
        raise SystemExit(f"rank {rank}: sort result failed verification (unsorted pairs: {unsorted})")

    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    job_ms = float(t.item())
    ms_per_step = job_ms / args.steps
    value = world * n / (ms_per_step * 1e-3) / 1e9

    # ---- roofline of the dominant kernel (K3 onesweep pass), timed live with CUDA events
    n_local = int(out.shape[0])
    rf_n = n if world == 1 else n_local
    pass_ms, hist_ms = [], []
    for _ in range(2):
        keys2 = keys[:rf_n] if world == 1 else out
        if world == 1:
            keys.copy_(pristine)
        tt = rs.sort_timed(keys2, K, workspace=ws, flags=rs.FLAG_NO_SKIP)
        pass_ms += [p for p in tt["pass_ms"]]
        hist_ms.append(tt["hist_ms"])
    avg_pass = sum(pass_ms) / len(pass_ms)
    peak, peak_src = peaks()
    achieved = 2.0 * rf_n * KEY_BYTES / (avg_pass * 1e-3) / 1e9
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "r01_traffic.json")) as fh:
            tj = json.load(fh)
        if int(tj.get("n", 0)) == rf_n and tj.get("kind") == KIND_NAME:
            traffic = int(tj["dram_bytes_read"]) + int(tj["dram_bytes_write"])
    except Exception:  # noqa: BLE001
        pass

    # ---- the reference's timer (main.rs:32-34) also covers the allocation and release of the temp buffer
    # (mod.rs:71-83, dropped at :176): the same sort with cudaMalloc / cudaFree of the workspace inside, wall clock
    alloc_inside = None
    if world == 1 and not args.no_u32:
        ws_bytes = int(ws.numel())
        del ws
        torch.cuda.empty_cache()
        ta = []
        for _ in range(3):
            keys.copy_(pristine)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            w = torch.empty(rs.temp_bytes(n, K), dtype=torch.uint8, device=dev)  # cache is empty: a real cudaMalloc
            rs.radix_sort(keys, K, workspace=w)
            torch.cuda.synchronize()
            del w
            torch.cuda.empty_cache()  # cudaFree
            ta.append(time.perf_counter() - t0)
        alloc_inside = {"value": round(n / min(ta) / 1e9, 3), "unit": "Gkeys/s", "ms_per_step": round(min(ta) * 1e3, 3),
                        "note": "device-resident keys; cudaMalloc + cudaFree of the workspace inside the timed region, host wall clock, best of 3"}
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)

    # ---- the metric names u32 too: the same workload shape with 4-byte keys (N = 1 only), reported beside the headline
    u32_line = None
    if world == 1 and not args.no_u32:
        K4 = rs.Kind.U32
        p4 = distr.empty_keys(n, K4, dev)
        distr.generate(p4, K4, distr.Distr.UNIFORM, 0, seed=SEED)
        k4 = torch.empty_like(p4)
        fp4 = verify.fingerprint(p4, K4)
        t4 = 0.0
        for i in range(2 + args.steps):
            k4.copy_(p4)
            torch.cuda.synchronize()
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record()
            rs.radix_sort(k4, K4, workspace=ws)
            a1.record()
            torch.cuda.synchronize()
            if i >= 2:
                t4 += a0.elapsed_time(a1)
        if verify.count_unsorted(k4, K4) != 0 or verify.fingerprint(k4, K4) != fp4:
            raise SystemExit("u32 sort failed verification")
        k4.copy_(p4)
        tt4 = rs.sort_timed(k4, K4, workspace=ws, flags=rs.FLAG_NO_SKIP)
        ap4 = sum(tt4["pass_ms"]) / len(tt4["pass_ms"])
        u32_line = {"value": round(n / (t4 / args.steps * 1e-3) / 1e9, 3), "unit": "Gkeys/s", "ms_per_step": round(t4 / args.steps, 4),
                    "workload": f"uniform random u32 keys, {n} elements, single B200", "pass_GBps": round(2.0 * n * 4 / (ap4 * 1e-3) / 1e9, 1),
                    "pass_frac": round(2.0 * n * 4 / (ap4 * 1e-3) / 1e9 / peak, 4)}
        del p4, k4

    # ---- e2e: host buffers through the drop-in C-ABI call, copies inside the timed region
    e2e_steps = 0 if args.no_e2e else max(1, min(args.steps, 5))
    host = torch.empty(1 if args.no_e2e else n if world == 1 else int(n * 1.25) + 1024, dtype=torch.uint64, pin_memory=True)
    e2e_total = 0.0
    h2d = d2h = n * KEY_BYTES
    for i in range(e2e_steps + 1 if e2e_steps else 0):
        host[:n].copy_(pristine)  # untimed: a fresh unsorted HOST input
        barrier()
        t0 = time.perf_counter()
        if world == 1:
            rc = _ffi.lib().b200sort_keys_host(host.data_ptr(), n, int(K))
            _ffi.check(rc)
        else:
            keys.copy_(host[:n], non_blocking=True)
            res = mgpu.sort_sharded(keys, K, workspace=ws, send_buffer=sendbuf, exchange=exchange, overlap=overlap,
                                    workspace2=ws2)
            host[: res.n_out].copy_(res.keys, non_blocking=True)
            torch.cuda.synchronize()
            d2h = res.n_out * KEY_BYTES
        dt = time.perf_counter() - t0
        if i > 0:  # first one is the warm-up (device pool allocation)
            e2e_total += dt
    te = torch.tensor([e2e_total], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = world * n / (float(te.item()) / e2e_steps) / 1e9 if e2e_steps else None
    if world == 1 and e2e_steps:
        hs = host[:n]
        assert bool((hs[1:].view(torch.int64) ^ (-2**63) >= hs[:-1].view(torch.int64) ^ (-2**63)).all()), "e2e result not sorted"

    # the same call on PAGEABLE host memory (what a Rust Vec<T> is): the library stages it through
    # its own pinned bounce buffers with copy threads.  Reported beside the contract's pinned number.
    e2e_pageable = None
    if world == 1 and e2e_steps:
        pristine_host = pristine.cpu().numpy()
        pageable = np.empty(n, dtype=np.uint64)
        tp = 0.0
        for i in range(3):  # the first one is the warm-up (bounce buffers, copy threads)
            pageable[...] = pristine_host
            t0 = time.perf_counter()
            _ffi.check(_ffi.lib().b200sort_keys_host(pageable.ctypes.data, n, int(K)))
            if i > 0:
                tp += time.perf_counter() - t0
        assert bool(np.all(pageable[1:] >= pageable[:-1])), "pageable e2e result not sorted"
        e2e_pageable = {"value": round(n / (tp / 2) / 1e9, 3), "unit": "Gkeys/s",
                        "note": "b200sort_keys_host on an unpinned numpy array (a Rust Vec is pageable too): staged through the "
                                "library's pinned bounce buffers by its copy threads"}
        del pageable, pristine_host

    if rank != 0:
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return 0

    cpu = None
    if world == 1 and not args.no_cpu:
        cores, res = cpu_baseline(CPU_SAMPLE_KEYS)
        cpu = {"value": round(res["radix_sort"], 4), "unit": "Gkeys/s", "cores": cores, "kind": "port",
               "sample": (f"{CPU_SAMPLE_KEYS} uniform u64 keys (1/10 of the workload), one run each: C restatement of "
                          f"RadixSort::radix_sort (mod.rs:61-176) = {res['radix_sort']:.4f} Gkeys/s; of the rayon variant "
                          f"radix_sort4 (mod.rs:405-466) = {res['radix_sort4']:.4f} Gkeys/s; {cores} host threads")}

    # NVLink roofline of the exchange (rank 0's phase time): every rank sends (world-1)/world of its keys
    nvlink = None
    xms = phases.get("exchange") or phases.get("all_to_all")
    if world > 1 and xms:
        sent = (world - 1) / world * n * KEY_BYTES
        if overlap:  # the timed "exchange" phase is the first of the two pieces: half of the bytes
            sent /= 2
        nvlink = {"sent_bytes_per_gpu": int(sent), "exchange_ms": round(xms, 3), "GBps_per_direction": round(sent / (xms * 1e-3) / 1e9, 1),
                  "frac_of_900": round(sent / (xms * 1e-3) / 1e9 / 900.0, 3)}
    line = {
        "metric": "sorted Gkeys/s (u64)", "value": round(value, 3), "unit": "Gkeys/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_per_step, 4),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": ("uniform random u64 keys, 1B elements, single B200 (BASELINE.json configs[1])" if world == 1 else
                                f"uniform random u64 keys, {n} per GPU x {world} GPUs, sharded; MSD splitters + exchange + local LSD "
                                "(weak-scaled form of BASELINE.json configs[4])"),
                   "exchange": exchange_note, "phases_ms_rank0": {k_: round(v_, 3) for k_, v_ in phases.items()},
                   "nvlink": nvlink,
                   "keys_per_gpu": n, "seed": SEED, "digits": DIGITS, "digit_bits": 8,
                   "l2": "inputs (8 GB/GPU) exceed the 126 MB L2; no flush between steps",
                   "e2e_steps": e2e_steps, "verified": "sorted + multiset fingerprint == input (bit-identical to the reference's output)"},
        "e2e": {"value": round(e2e_value, 3) if e2e_value is not None else None, "unit": "Gkeys/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "hbm", "kernel": "onesweep pass (K3), one 8-bit digit", "achieved": round(achieved, 1), "peak": peak,
                     "unit": "GB/s", "frac": round(achieved / peak, 4), "traffic": traffic,
                     "peak_source": peak_src, "bytes_per_launch": 2 * rf_n * KEY_BYTES, "avg_launch_ms": round(avg_pass, 4),
                     "hist_kernel_GBps": round(rf_n * KEY_BYTES / (sum(hist_ms) / len(hist_ms) * 1e-3) / 1e9, 1),
                     "whole_sort_GBps": round(KEY_BYTES * (1 + 2 * DIGITS) * world * n / (ms_per_step * 1e-3) / 1e9 / world, 1)},
    }
    if u32_line is not None:
        line["u32"] = u32_line
    if alloc_inside is not None:
        line["alloc_inside"] = alloc_inside
    if e2e_pageable is not None:
        line["e2e_pageable"] = e2e_pageable
    if cpu is not None:
        line["cpu_baseline"] = cpu
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--keys", dest="n", type=int, default=N_KEYS, help="keys per GPU (default: the BASELINE.json workload)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-u32", action="store_true", help="skip the secondary u32 measurement")
    ap.add_argument("--no-e2e", action="store_true",
                    help="skip the host-buffer leg (evidence runs at sizes whose pinned host copies would not fit the box)")
    ap.add_argument("--overlap", default="off", choices=["on", "off"],
                    help="N > 1, fused exchange: send the buckets in two pieces and sort the first while the second travels")
    ap.add_argument("--exchange", default="auto", choices=["auto", "fused", "nccl"],
                    help="N > 1: fused K7 peer stores, NCCL all-to-all, or fused with fallback (default)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
