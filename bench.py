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

e2e = the same metric through the drop-in C-ABI call on ONE pinned HOST slice,
H2D copy, sort, D2H copy all inside the timed region: b200sort_keys_host at
N = 1, b200sort_mgpu_keys_host (one process, all N GPUs; rank 0 makes the call
while the other ranks wait on a host-side barrier) at N > 1.

The other BASELINE.json configurations (C1 10 M u32, C3 1 B f32/f64/i32/i64,
C4 2 B skewed / few-unique / presorted u32, the demo's record shapes, and C5
16 B keys on 8 GPUs) are timed after the headline and reported inside
`roofline` (`u32`, `other_configs`, `c5_16B_keys`), where the driver keeps them.

--impl reference times the reference's CPU algorithm (the C restatement under
oracle/, since no Rust toolchain exists on these boxes) on the host cores, on the
same configuration: N = 1 sorts the whole 1 B-key workload every step.
"""
from __future__ import annotations

import argparse
import ctypes
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
C5_KEYS_PER_GPU = 2_000_000_000  # BASELINE.json configs[4]: 16 B keys over 8 GPUs


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def make_config(world: int, n: int):
    """The `config` object of the JSON line: the same for both arms (the driver compares them)."""
    if world == 1:
        workload = (f"uniform random u64 keys, {'1B' if n == N_KEYS else n} elements, single B200"
                    + (" (BASELINE.json configs[1])" if n == N_KEYS else ""))
    else:
        workload = (f"uniform random u64 keys, {n} per GPU x {world} GPUs, sharded; MSD splitters + exchange + local LSD "
                    "(weak-scaled form of BASELINE.json configs[4])")
    return {"workload": workload, "keys_per_gpu": n, "n_gpus": world, "seed": SEED, "digits": DIGITS, "digit_bits": 8,
            "l2": "inputs (8 GB/GPU) exceed the 126 MB L2; no flush between steps"}


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

    def mark(self):
        """rows from here on belong to the timed region"""
        self.first = len(self.rows)

    def stop(self):
        if getattr(self, "first", 0) and len(self.rows) > self.first:
            self.rows = self.rows[self.first:]
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


# ---------------------------------------------------------------------------------------
# CPU arms (oracle/: the C restatement of the reference's sorts; test infrastructure, timed here
# as the stated baseline -- never part of the GPU path)
# ---------------------------------------------------------------------------------------
def cpu_sort_seconds(fn, src, work, kind: str, cores: int) -> float:
    import numpy as np
    np.copyto(work, src)
    t0 = time.perf_counter()
    fn(work, kind, cores)
    return time.perf_counter() - t0


def cpu_baseline(n: int):
    """One run each, on all host cores, of the oracle's restatement of the reference's production
    sort (mod.rs:61-176, native threads + 96-element write buffers) and of its rayon variant
    radix_sort4 (mod.rs:405-466), on n keys of the headline workload."""
    import numpy as np
    from oracle import cpu_oracle
    cores = cpu_oracle.available_parallelism()
    src = cpu_oracle.generate_uniform(n, "u64", SEED)
    work = np.empty_like(src)
    out = {}
    for name, fn in (("radix_sort", cpu_oracle.sort_mt), ("radix_sort4", cpu_oracle.sort4_mt)):
        dt = cpu_sort_seconds(fn, src, work, "u64", cores)
        assert bool(np.all(work[:-1] <= work[1:]))
        out[name] = n / dt / 1e9
    return cores, out


def cpu_records_baseline(pairs: int):
    """The demo's smallest (u32,u32) case (main.rs:104: 0.5 GB) through the oracle's production sort."""
    import numpy as np
    from oracle import cpu_oracle
    cores = cpu_oracle.available_parallelism()
    src = np.zeros((pairs, 2), dtype=np.uint32)                # KeyUniform: (Standard, 0) -- distr.rs:42-46
    src[:, 0] = cpu_oracle.generate_uniform(pairs, "u32", SEED)
    work = np.empty_like(src)
    dt = cpu_sort_seconds(cpu_oracle.sort_mt, src, work, "pair_u32", cores)
    assert bool(np.all(work[:-1, 0] <= work[1:, 0]))
    return cores, pairs / dt / 1e9


def run_reference(args):
    """--impl reference: the reference's CPU path on this box's host cores, same configuration."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import numpy as np
    from oracle import cpu_oracle
    cores = cpu_oracle.available_parallelism()
    world = args.gpus
    n = args.n  # N = 1: the whole workload; N > 1: one GPU's share of it (a host cannot hold N x 8 GB twice)
    src = cpu_oracle.generate_uniform(n, "u64", SEED)
    work = np.empty_like(src)
    for _ in range(args.warmup):
        cpu_sort_seconds(cpu_oracle.sort_mt, src, work, "u64", cores)
    total = 0.0
    for _ in range(args.steps):
        total += cpu_sort_seconds(cpu_oracle.sort_mt, src, work, "u64", cores)
    assert bool(np.all(work[:-1] <= work[1:]))
    rayon = n / cpu_sort_seconds(cpu_oracle.sort4_mt, src, work, "u64", cores) / 1e9
    ms = total / args.steps * 1e3
    value = n / (ms * 1e-3) / 1e9
    sample = ((f"the whole workload: {n} uniform u64 keys per step" if world == 1 else
               f"{n} uniform u64 keys per step = one GPU's share of the {world} x {n}-key workload (per-key throughput; "
               "a host cannot hold the whole of it twice)")
              + f"; oracle_sort_mt = C restatement of RadixSort::radix_sort (mod.rs:61-176), {cores} threads, temp allocation "
                f"inside the timer as main.rs:32-34; the rayon variant radix_sort4 (mod.rs:405-466), one run: {rayon:.4f} Gkeys/s; "
                "the Rust crate itself cannot be built here (no cargo/rustc)")
    line = {
        "impl": "reference", "metric": "sorted Gkeys/s (u64)", "value": round(value, 4), "unit": "Gkeys/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": make_config(world, n),
        "cpu_baseline": {"value": round(value, 4), "unit": "Gkeys/s", "cores": cores, "kind": "port", "sample": sample,
                         "radix_sort4_gkeys_s": round(rayon, 4)},
        "e2e": {"value": round(value, 4), "unit": "Gkeys/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


# ---------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------
def timed_sorts(rs, torch, keys, pristine, K, ws, steps: int, warm: int = 1):
    """average device ms of `steps` sorts of `pristine` (restored outside the timed region)"""
    total = 0.0
    for i in range(warm + steps):
        keys.copy_(pristine)
        torch.cuda.synchronize()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record()
        rs.radix_sort(keys, K, workspace=ws)
        a1.record()
        torch.cuda.synchronize()
        if i >= warm:
            total += a0.elapsed_time(a1)
    return total / steps


def other_configs(rs, torch, np, distr, verify, _ffi, dev, buf_a, buf_b, ws, peak, steps: int, with_cpu: bool, note_error, failed):
    """BASELINE.json configs C1, C3, C4 and the demo's record shapes at their stated sizes, on one GPU.
    buf_a / buf_b: two 8 GB device buffers (uint8), reused for every case."""
    out = {}
    steps = max(1, min(steps, 3))

    def views(kind, n):
        kb = rs.key_bytes(kind)
        dt = distr._TORCH_DTYPE[rs.Kind(kind)]
        a = buf_a[: n * kb].view(getattr(torch, dt))
        b = buf_b[: n * kb].view(getattr(torch, dt))
        return a, b, kb

    def run_case(kind, n, dist_id, param, seed, label):
        K = rs.Kind[kind.upper()]
        pristine, keys, kb = views(K, n)
        distr.generate(pristine, K, dist_id, param, seed=seed)
        fp = verify.fingerprint(pristine, K)
        ms = timed_sorts(rs, torch, keys, pristine, K, ws, steps)
        ok = verify.count_unsorted(keys, K) == 0 and verify.fingerprint(keys, K) == fp
        if not ok:
            raise SystemExit(f"{label}: sort failed verification")
        keys.copy_(pristine)
        tt = rs.sort_timed(keys, K, workspace=ws)
        executed = [p for p in tt["pass_ms"] if p > 0.25 * max(tt["pass_ms"])]
        bytes_moved = n * kb * (1 + 2 * len(executed))
        return {"n": n, "kind": kind, "ms": round(ms, 4), "gkeys_s": round(n / ms / 1e6, 3),
                "executed_passes": len(executed), "digits": len(tt["pass_ms"]),
                "executed_bytes": bytes_moved, "executed_GBps": round(bytes_moved / ms / 1e6, 1),
                "pass_frac_of_peak": round(2.0 * n * kb / (sum(executed) / len(executed)) / 1e6 / peak, 4) if executed else None,
                "verified": True}

    def case(name, *a):
        """one configuration; a failure is recorded under its own name and the others still run"""
        try:
            out[name] = run_case(*a)
        except BaseException as e:  # noqa: BLE001
            note_error(f"other_configs/{name}", e)
            out[name] = {"error": f"{type(e).__name__}: {e}"[:200], "verified": False}
            if isinstance(e, (SystemExit, AssertionError)):
                failed.append(name)

    D = distr.Distr
    # C1: 10 M uniform u32 (L2-resident: 40 MB + 40 MB temp; launch-latency bound, no HBM fraction claimed)
    try:
        c1 = run_case("u32", 10_000_000, D.UNIFORM, 0.0, 0xB2000001, "C1")
        c1.pop("pass_frac_of_peak")
        c1["note"] = "device-resident; 80 MB working set sits in the 126 MB L2"
        host = np.asarray(distr.uniform_host(10_000_000, rs.Kind.U32, seed=0xB2000001))
        work = host.copy()
        best = None
        for _ in range(4):
            np.copyto(work, host)
            t0 = time.perf_counter()
            _ffi.check(_ffi.lib().b200sort_keys_host(work.ctypes.data, work.size, int(rs.Kind.U32)))
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        assert bool(np.all(work[1:] >= work[:-1]))
        c1["host_call_ms"] = round(best * 1e3, 3)
        c1["host_call_note"] = "b200sort_keys_host on a pageable numpy array (what a Rust Vec is), copies inside, best of 3 after a warm-up"
        out["C1_10M_u32"] = c1
    except BaseException as e:  # noqa: BLE001
        note_error("other_configs/C1_10M_u32", e)
        out["C1_10M_u32"] = {"error": f"{type(e).__name__}: {e}"[:200], "verified": False}
        if isinstance(e, (SystemExit, AssertionError)):
            failed.append("C1_10M_u32")
    # C3: 1 B keys of the transform kinds (radix_digits.rs:55-124)
    n3 = 1_000_000_000
    case("C3_1B_f32_test_shaped", "f32", n3, D.FLOAT_TEST, 0.0, 0xB2000003, "C3 f32 test-shaped")
    case("C3_1B_f32_raw_bits", "f32", n3, D.UNIFORM, 0.0, 0xB2000003, "C3 f32 raw")
    case("C3_1B_f64_raw_bits", "f64", n3, D.UNIFORM, 0.0, 0xB2000003, "C3 f64 raw")
    case("C3_1B_i32", "i32", n3, D.UNIFORM, 0.0, 0xB2000003, "C3 i32")
    case("C3_1B_i64", "i64", n3, D.UNIFORM, 0.0, 0xB2000003, "C3 i64")
    # C4: 2 B u32, skewed / few-unique / presorted (distr.rs families; presorted is the builder's)
    n4 = 2_000_000_000
    case("C4_2B_u32_zipf_1.0", "u32", n4, D.ZIPF, 1.0, 0xB2000004, "C4 zipf")
    case("C4_2B_u32_step_16", "u32", n4, D.STEP, 16, 0xB2000004, "C4 step 16")
    case("C4_2B_u32_ascending", "u32", n4, D.ASCENDING, 0.0, 0xB2000004, "C4 ascending")
    case("C4_2B_u32_geometric_2^-10", "u32", n4, D.GEOMETRIC, 2.0 ** -10, 0xB2000004, "C4 geometric")
    # the demo's record shapes (main.rs:104-125): (u32,u32) and (u64,u64) pairs, payload 0 (distr.rs:42-52), 4.0 GB
    for name, kind, pairs in (("records_4GB_u32_u32", "u32", 500_000_000), ("records_4GB_u64_u64", "u64", 250_000_000)):
        try:
            K = rs.Kind[kind.upper()]
            kb = rs.key_bytes(K)
            dt = getattr(torch, distr._TORCH_DTYPE[K])
            pristine = buf_a[: pairs * 2 * kb].view(dt).view(pairs, 2)
            recs = buf_b[: pairs * 2 * kb].view(dt).view(pairs, 2)
            col = buf_b[4_000_000_000: 4_000_000_000 + pairs * kb].view(dt)
            distr.generate(col, K, D.UNIFORM, 0.0, seed=0xB2000005)
            pristine.zero_()
            pristine[:, 0] = col
            total = 0.0
            for i in range(1 + steps):
                recs.copy_(pristine)
                torch.cuda.synchronize()
                a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a0.record()
                rs.radix_sort_records(recs, K, 0, workspace=ws)
                a1.record()
                torch.cuda.synchronize()
                if i >= 1:
                    total += a0.elapsed_time(a1)
            ms = total / steps
            keycol = recs[:, 0].contiguous()
            if verify.count_unsorted(keycol, K) != 0 or verify.fingerprint(keycol, K) != verify.fingerprint(col, K):
                raise SystemExit(f"{name}: record sort failed verification")
            out[name] = {"records": pairs, "record_bytes": 2 * kb, "ms": round(ms, 4), "grecords_s": round(pairs / ms / 1e6, 3),
                         "GBps": round(pairs * 2 * kb * (1 + 2 * kb) / ms / 1e6, 1),
                         "reference_shape": "main.rs:104-125 (KeyUniform pairs, payload 0), largest size", "verified": "keys sorted + key multiset"}
        except BaseException as e:  # noqa: BLE001
            note_error(f"other_configs/{name}", e)
            out[name] = {"error": f"{type(e).__name__}: {e}"[:200], "verified": False}
            if isinstance(e, (SystemExit, AssertionError)):
                failed.append(name)
    if with_cpu:
        try:
            cores, g = cpu_records_baseline(62_500_000)
            out["records_cpu_baseline"] = {"grecords_s": round(g, 4), "cores": cores, "kind": "port",
                                           "sample": "62.5 M (u32,u32) pairs = the demo's 0.5 GB case (main.rs:104), oracle_sort_mt on pair_u32, one run"}
        except BaseException as e:  # noqa: BLE001
            note_error("other_configs/records_cpu_baseline", e)
    return out


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
    host_group = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29531")
        # stdout carries the ONE JSON line: NCCL's version banner (NCCL_DEBUG=VERSION on some boxes) goes to stderr
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
        # host-side barrier (gloo): while rank 0 drives all GPUs from one process (the drop-in call, the
        # communicator parity case) the other ranks must not have a spinning NCCL kernel on their GPU
        host_group = dist.new_group(backend="gloo")

    def host_barrier():
        torch.cuda.synchronize()
        if host_group is not None:
            dist.barrier(group=host_group)

    # Nothing after the timed steps may cost the record: every auxiliary leg (parity pre-step, configs[4], alloc-inside,
    # u32, the other configurations, e2e, cpu_baseline) runs guarded, its failure lands in the line's "errors" list with
    # the traceback on stderr, the ONE JSON line is printed all the same, and the exit code is non-zero afterwards when a
    # parity / verification leg was among the failures.
    errors, failed = [], []

    def note_error(leg, e):
        import traceback
        sys.stderr.write(f"[bench.py rank {rank}] leg '{leg}' failed:\n{''.join(traceback.format_exception(type(e), e, e.__traceback__))}\n")
        sys.stderr.flush()
        errors.append({"leg": leg, "rank": rank, "msg": f"{type(e).__name__}: {e}"[:300]})

    K = rs.Kind.U64
    n = args.n
    c5 = world >= 8 and not args.no_extra and n == N_KEYS  # configs[4] itself, as an extra labelled point
    cap_keys = int((C5_KEYS_PER_GPU if c5 else n) * 1.25) + 1024

    # ---- N > 1: parity of the multi-GPU paths, untimed, before anything is measured
    mgpu_parity = None
    if world > 1 and not args.no_parity:
        from tests import _mgpu_worker
        perr = 0
        try:
            _mgpu_worker.run_cases(dist, dev, rank, world)        # rank-per-process: fused exchange, overlap, NCCL fallback
        except BaseException as e:  # noqa: BLE001  (recorded; the line is still printed and the exit code says so)
            perr = 1
            note_error("mgpu_parity/run_cases", e)
        host_barrier()
        if rank == 0:
            try:
                _mgpu_worker.communicator_case(list(range(world)))  # one process over all GPUs (b200sort_mgpu_*)
            except BaseException as e:  # noqa: BLE001
                perr = 1
                note_error("mgpu_parity/communicator", e)
        host_barrier()
        pf = torch.tensor([perr], dtype=torch.int32, device=dev)
        dist.all_reduce(pf, op=dist.ReduceOp.MAX)
        mgpu_parity = "ok" if int(pf.item()) == 0 else "FAILED"
        if mgpu_parity != "ok":
            failed.append("mgpu_parity")

    pristine = distr.empty_keys(n, K, dev)
    distr.generate(pristine, K, distr.Distr.UNIFORM, 0, seed=SEED, index_base=rank * n)
    keys = torch.empty_like(pristine)
    ws = torch.empty(rs.temp_bytes(cap_keys, K), dtype=torch.uint8, device=dev)
    fp_in = verify.fingerprint(pristine, K)
    # N > 1: fused K7 + exchange over peer memory (torch symmetric memory) unless
    # it cannot be set up here; then the NCCL all-to-all path, and the line says so
    exchange, exchange_note, sendbuf = None, "single GPU", None
    if world > 1:
        exchange_note = "nccl all_to_all_single"
        if args.exchange in ("auto", "fused"):
            try:
                exchange = mgpu.PeerExchange(cap_keys, KEY_BYTES, dev)
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
    claim = exchange is not None and args.claim == "on" and not overlap
    if claim:
        exchange_note += "; count-free: K7 tiles claim room in the receive buffers through cursors on the destination GPUs"
    if overlap:
        exchange_note += "; exchange in two pieces, the second overlapped with the local sort of the first"

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def sharded(k_in, timings=None):
        return mgpu.sort_sharded(k_in, K, workspace=ws, send_buffer=sendbuf, exchange=exchange, overlap=overlap,
                                 workspace2=ws2, timings=timings, claim=claim)

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
            out = sharded(keys).keys
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1), out

    def verify_result(out, fp_before, what):
        """sorted + same multiset (count, sum mix64, xor mix64, sum key) <=> bit-identical to the reference's
        output; N > 1: the per-rank fingerprints are combined over the ranks, and shard boundaries must ascend"""
        unsorted = verify.count_unsorted(out, K)
        fp_out = verify.fingerprint(out, K)
        ok = unsorted == 0
        if world == 1:
            ok = ok and fp_out == fp_before
        else:
            s64 = lambda v: v - 2**64 if v >= 2**63 else v  # noqa: E731  (bit pattern of a u64 in an int64 tensor)
            t = torch.tensor([s64(v) for v in (*fp_before, *fp_out)], dtype=torch.int64, device=dev)
            both = [torch.zeros_like(t) for _ in range(world)]
            dist.all_gather(both, t)
            tot = [0] * 8
            for b in both:
                v = [int(x) % 2**64 for x in b]
                for i in range(8):
                    tot[i] = (tot[i] ^ v[i]) if i in (2, 6) else (tot[i] + v[i]) % 2**64
            ok = ok and tot[0:4] == tot[4:8]
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
                if int(b[4]) == 0:
                    continue
                if last is not None and int(e[0]) < last:
                    ok = False
                last = int(e[1])
        if world > 1:  # one verdict for all ranks, so that nobody is left waiting in the next collective
            okt = torch.tensor([1 if ok else 0], dtype=torch.int32, device=dev)
            dist.all_reduce(okt, op=dist.ReduceOp.MIN)
            if ok and int(okt.item()) == 0:
                raise SystemExit(f"rank {rank}: {what}: another rank's part of the result failed verification")
        if not ok:
            raise SystemExit(f"rank {rank}: {what}: sort result failed verification (unsorted pairs: {unsorted})")

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()  # before the warm-up: nvidia-smi takes a while to print its first row
    for _ in range(max(args.warmup, 0)):
        one_step()
    launches0 = rs.launch_count()
    barrier()
    sampler.mark()
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
        out = sharded(keys, timings=phases).keys
        barrier()

    # correctness of the last step
    verify_result(out, fp_in, "headline")

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
    for name in ("r03_traffic.json", "r02_traffic.json", "r01_traffic.json"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as fh:
                tj = json.load(fh)
            if int(tj.get("n", 0)) == rf_n and tj.get("kind") == KIND_NAME:
                traffic = int(tj["dram_bytes_read"]) + int(tj["dram_bytes_write"])
                break
        except Exception:  # noqa: BLE001
            pass

    # ---- configs[4] as stated (16 B u64 keys over 8 GPUs = 2 B per GPU), an extra labelled point at N >= 8
    c5_line = None
    if c5:
        try:
            del pristine, keys, out
            torch.cuda.empty_cache()
            n5 = C5_KEYS_PER_GPU
            k5 = distr.empty_keys(n5, K, dev)
            t5 = 0.0
            c5_steps = 2
            for i in range(1 + c5_steps):
                distr.generate(k5, K, distr.Distr.UNIFORM, 0, seed=SEED + 5, index_base=rank * n5)  # untimed: fresh unsorted input
                if i == 0:
                    fp5 = verify.fingerprint(k5, K)
                barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                o5 = sharded(k5).keys
                e1.record()
                torch.cuda.synchronize()
                if i >= 1:
                    t5 += e0.elapsed_time(e1)
            verify_result(o5, fp5, "configs[4]")
            tt5 = torch.tensor([t5], dtype=torch.float64, device=dev)
            dist.all_reduce(tt5, op=dist.ReduceOp.MAX)
            ms5 = float(tt5.item()) / c5_steps
            c5_line = {"workload": f"BASELINE.json configs[4]: {world * n5} uniform u64 keys sharded over {world} B200 ({n5} per GPU)",
                       "keys_per_gpu": n5, "ms_per_step": round(ms5, 3), "value": round(world * n5 / ms5 / 1e6, 3), "unit": "Gkeys/s",
                       "steps": c5_steps, "verified": "sorted + multiset fingerprint over all ranks + ascending shard boundaries"}
            del k5, o5
        except BaseException as e:  # noqa: BLE001
            note_error("configs[4]", e)
            if isinstance(e, (SystemExit, AssertionError)):
                failed.append("configs[4]")
        k5 = o5 = None
        torch.cuda.empty_cache()
        pristine = distr.empty_keys(n, K, dev)
        distr.generate(pristine, K, distr.Distr.UNIFORM, 0, seed=SEED, index_base=rank * n)
        keys = torch.empty_like(pristine)

    # ---- the reference's timer (main.rs:32-34) also covers the allocation and release of the temp buffer
    # (mod.rs:71-83, dropped at :176): the same sort with cudaMalloc / cudaFree of the workspace inside, wall clock
    alloc_inside = None
    if world == 1 and not args.no_u32:
        ws_bytes = int(ws.numel())
        try:
            ws = None
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
            # the same convention through b200sort_keys_device_pooled: the workspace comes from the library's stream-ordered pool
            tp_ = []
            st_ = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
            for i in range(4):  # the first call grows the pool
                keys.copy_(pristine)
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                _ffi.check(_ffi.lib().b200sort_keys_device_pooled(keys.data_ptr(), n, int(K), st_))
                torch.cuda.synchronize()
                if i > 0:
                    tp_.append(time.perf_counter() - t0)
            if verify.count_unsorted(keys, K) != 0 or verify.fingerprint(keys, K) != fp_in:
                raise SystemExit("pooled-workspace sort failed verification")
            _ffi.lib().b200sort_release_cached_memory()
            alloc_inside["pooled"] = {"value": round(n / min(tp_) / 1e9, 3), "unit": "Gkeys/s", "ms_per_step": round(min(tp_) * 1e3, 3),
                                      "note": "b200sort_keys_device_pooled: workspace taken and returned inside the call from the library's "
                                              "stream-ordered pool (cudaMallocFromPoolAsync / cudaFreeAsync), host wall clock, best of 3 after the first call"}
            ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        except BaseException as e:  # noqa: BLE001
            note_error("alloc_inside", e)
        if ws is None:
            torch.cuda.empty_cache()
            ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)

    # ---- the metric names u32 too: the same workload shape with 4-byte keys (N = 1 only), reported beside the headline
    u32_line = None
    if world == 1 and not args.no_u32:
        try:
            K4 = rs.Kind.U32
            p4 = distr.empty_keys(n, K4, dev)
            distr.generate(p4, K4, distr.Distr.UNIFORM, 0, seed=SEED)
            k4 = torch.empty_like(p4)
            fp4 = verify.fingerprint(p4, K4)
            ms4 = timed_sorts(rs, torch, k4, p4, K4, ws, args.steps, warm=2)
            if verify.count_unsorted(k4, K4) != 0 or verify.fingerprint(k4, K4) != fp4:
                raise SystemExit("u32 sort failed verification")
            k4.copy_(p4)
            tt4 = rs.sort_timed(k4, K4, workspace=ws, flags=rs.FLAG_NO_SKIP)
            ap4 = sum(tt4["pass_ms"]) / len(tt4["pass_ms"])
            u32_line = {"value": round(n / (ms4 * 1e-3) / 1e9, 3), "unit": "Gkeys/s", "ms_per_step": round(ms4, 4),
                        "workload": f"uniform random u32 keys, {n} elements, single B200", "pass_GBps": round(2.0 * n * 4 / (ap4 * 1e-3) / 1e9, 1),
                        "pass_frac": round(2.0 * n * 4 / (ap4 * 1e-3) / 1e9 / peak, 4), "avg_launch_ms": round(ap4, 4),
                        "whole_sort_GBps": round(36.0 * n / (ms4 * 1e-3) / 1e9, 1)}
            del p4, k4
        except BaseException as e:  # noqa: BLE001
            note_error("u32", e)
            if isinstance(e, (SystemExit, AssertionError)):
                failed.append("u32")

    # ---- the other configurations of BASELINE.json, at their stated sizes (N = 1)
    others = None
    if world == 1 and not args.no_extra and n == N_KEYS:
        try:
            buf_a = pristine.view(torch.uint8)
            buf_b = keys.view(torch.uint8)
            others = other_configs(rs, torch, np, distr, verify, _ffi, dev, buf_a, buf_b, ws, peak, args.steps, not args.no_cpu, note_error, failed)
            distr.generate(pristine, K, distr.Distr.UNIFORM, 0, seed=SEED, index_base=rank * n)  # the buffers were reused
        except BaseException as e:  # noqa: BLE001
            note_error("other_configs", e)

    # ---- e2e: ONE pinned host slice through the drop-in C-ABI call, copies inside the timed region
    e2e_steps = 0 if args.no_e2e else max(1, min(args.steps, 5))
    e2e_value, e2e_note, h2d, d2h = None, None, n * KEY_BYTES, n * KEY_BYTES
    e2e_pageable = None
    if e2e_steps and world == 1:
        try:
            host = torch.empty(n, dtype=torch.uint64, pin_memory=True)
            e2e_total = 0.0
            for i in range(e2e_steps + 1):
                host.copy_(pristine)  # untimed: a fresh unsorted HOST input
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                _ffi.check(_ffi.lib().b200sort_keys_host(host.data_ptr(), n, int(K)))
                dt = time.perf_counter() - t0
                if i > 0:  # first one is the warm-up (device pool allocation)
                    e2e_total += dt
            e2e_value = n / (e2e_total / e2e_steps) / 1e9
            e2e_note = "b200sort_keys_host on one pinned host slice"
            assert bool((host[1:].view(torch.int64) ^ (-2**63) >= host[:-1].view(torch.int64) ^ (-2**63)).all()), "e2e result not sorted"
            del host
            # the same call on PAGEABLE host memory (what a Rust Vec<T> is): the library stages it through
            # its own pinned bounce buffers with copy threads.  Reported beside the contract's pinned number.
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
        except BaseException as e:  # noqa: BLE001
            note_error("e2e (b200sort_keys_host)", e)
            if isinstance(e, (SystemExit, AssertionError)):
                failed.append("e2e (b200sort_keys_host)")
    elif e2e_steps:
        # N > 1: the reference's caller is ONE process with ONE slice: rank 0 calls b200sort_mgpu_keys_host over
        # all N GPUs on a pinned slice of N x n keys; the other ranks free their buffers and wait on the host
        keys = ws = exchange = sendbuf = ws2 = None  # free this rank's buffers: rank 0 drives every GPU now
        out = None
        torch.cuda.empty_cache()
        host_barrier()
        res = [0.0, 0.0]
        if rank == 0:
            try:
                total_keys = world * n
                try:
                    host = torch.empty(total_keys, dtype=torch.uint64, pin_memory=True)
                except Exception:  # noqa: BLE001  (the box cannot pin N x 8 GB: one GPU's share, and the line says so)
                    total_keys = n
                    host = torch.empty(total_keys, dtype=torch.uint64, pin_memory=True)
                comm = mgpu.Communicator(list(range(world)))
                host_np = host.numpy()  # Communicator.sort_host takes a writable numpy array (same pinned memory)
                tot = 0.0
                for i in range(e2e_steps + 1):
                    for r in range(total_keys // n):  # untimed: a fresh unsorted HOST input, the ranks' shards side by side
                        distr.generate(pristine, K, distr.Distr.UNIFORM, 0, seed=SEED, index_base=r * n)
                        host[r * n:(r + 1) * n].copy_(pristine)
                    torch.cuda.synchronize()
                    t0 = time.perf_counter()
                    comm.sort_host(host_np, K)
                    dt = time.perf_counter() - t0
                    if i > 0:
                        tot += dt
                hs = host.view(torch.int64) ^ (-2**63)
                step = 1 << 28
                for lo in range(0, total_keys - 1, step):
                    hi = min(total_keys, lo + step + 1)
                    assert bool((hs[lo + 1:hi] >= hs[lo:hi - 1]).all()), "e2e result not sorted"
                comm.close()
                res = [float(total_keys), tot / e2e_steps]
                del host_np, host
            except BaseException as e:  # noqa: BLE001
                note_error("e2e (b200sort_mgpu_keys_host)", e)
                if isinstance(e, (SystemExit, AssertionError)):
                    failed.append("e2e (b200sort_mgpu_keys_host)")
        host_barrier()
        tr = torch.tensor(res, dtype=torch.float64, device=dev)
        dist.broadcast(tr, src=0)
        total_keys, sec = int(tr[0].item()), float(tr[1].item())
        e2e_value = total_keys / sec / 1e9 if sec > 0 else None  # None: rank 0's leg failed (see "errors")
        h2d = d2h = total_keys * KEY_BYTES
        e2e_note = (f"b200sort_mgpu_keys_host: one process (rank 0), one pinned host slice of {total_keys} keys over all {world} GPUs"
                    + ("" if total_keys == world * n else " (the box could not pin N x 8 GB: one GPU's share of the workload)"))

    if rank != 0:
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return 1 if failed else 0

    cpu = None
    if world == 1 and not args.no_cpu:
        try:
            ncpu = min(n, args.cpu_keys)
            cores, res = cpu_baseline(ncpu)
            cpu = {"value": round(res["radix_sort"], 4), "unit": "Gkeys/s", "cores": cores, "kind": "port",
                   "sample": (f"{ncpu} uniform u64 keys ({'the whole workload' if ncpu == n else f'{ncpu / n:.2f} of the workload'}), one run each: "
                              f"C restatement of RadixSort::radix_sort (mod.rs:61-176) = {res['radix_sort']:.4f} Gkeys/s; of the rayon variant "
                              f"radix_sort4 (mod.rs:405-466) = {res['radix_sort4']:.4f} Gkeys/s; {cores} host threads")}
        except BaseException as e:  # noqa: BLE001
            note_error("cpu_baseline", e)

    # NVLink roofline of the exchange (rank 0's phase time): every rank sends (world-1)/world of its keys
    nvlink = None
    xms = phases.get("exchange") or phases.get("all_to_all")
    if world > 1 and xms:
        sent = (world - 1) / world * n * KEY_BYTES
        if overlap:  # the timed "exchange" phase is the first of the two pieces: half of the bytes
            sent /= 2
        nvlink = {"sent_bytes_per_gpu": int(sent), "exchange_ms": round(xms, 3), "GBps_per_direction": round(sent / (xms * 1e-3) / 1e9, 1),
                  "frac_of_900": round(sent / (xms * 1e-3) / 1e9 / 900.0, 3), "frac_of_770_measured_peer_copy": round(sent / (xms * 1e-3) / 1e9 / 770.0, 3)}
    roofline = {"bound": "hbm", "kernel": "onesweep pass (K3), one 8-bit digit", "achieved": round(achieved, 1), "peak": peak,
                "unit": "GB/s", "frac": round(achieved / peak, 4), "traffic": traffic,
                "peak_source": peak_src, "bytes_per_launch": 2 * rf_n * KEY_BYTES, "avg_launch_ms": round(avg_pass, 4),
                "hist_kernel_GBps": round(rf_n * KEY_BYTES / (sum(hist_ms) / len(hist_ms) * 1e-3) / 1e9, 1),
                "whole_sort_GBps": round(KEY_BYTES * (1 + 2 * DIGITS) * world * n / (ms_per_step * 1e-3) / 1e9 / world, 1),
                "exchange": exchange_note, "phases_ms_rank0": {k_: round(v_, 3) for k_, v_ in phases.items()}, "nvlink": nvlink,
                "e2e_steps": e2e_steps, "e2e_call": e2e_note,
                "verified": "sorted + multiset fingerprint (count, sum, xor, sum) == input (bit-identical to the reference's output)"}
    if mgpu_parity is not None:
        roofline["mgpu_parity"] = mgpu_parity
    if u32_line is not None:
        roofline["u32"] = u32_line
    if others is not None:
        roofline["other_configs"] = others
    if c5_line is not None:
        roofline["c5_16B_keys"] = c5_line
    line = {
        "metric": "sorted Gkeys/s (u64)", "value": round(value, 3), "unit": "Gkeys/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_per_step, 4),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": make_config(world, n),
        "e2e": {"value": round(e2e_value, 3) if e2e_value is not None else None, "unit": "Gkeys/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roofline,
    }
    if mgpu_parity is not None:
        line["mgpu_parity"] = mgpu_parity
    if alloc_inside is not None:
        line["alloc_inside"] = alloc_inside
    if e2e_pageable is not None:
        line["e2e_pageable"] = e2e_pageable
    if cpu is not None:
        line["cpu_baseline"] = cpu
    line["errors"] = errors
    # the numbers BASELINE.json's metric names ("sorted Gkeys/s (u32, u64)") and the other configurations, compact and LAST:
    # whoever keeps only the tail of this line keeps these
    oc = others or {}
    g = lambda name, key: (oc.get(name) or {}).get(key)  # noqa: E731
    summary = {
        "u64_gkeys_s": round(value, 3), "u64_pass_frac": round(achieved / peak, 4),
        "u32_gkeys_s": (u32_line or {}).get("value"), "u32_pass_frac": (u32_line or {}).get("pass_frac"),
        "e2e_gkeys_s": round(e2e_value, 3) if e2e_value is not None else None,
        "e2e_pageable_gkeys_s": (e2e_pageable or {}).get("value"),
        "alloc_inside_gkeys_s": (alloc_inside or {}).get("value"),
        "alloc_inside_pooled_gkeys_s": ((alloc_inside or {}).get("pooled") or {}).get("value"),
        "c1_ms": g("C1_10M_u32", "ms"), "c1_host_call_ms": g("C1_10M_u32", "host_call_ms"),
        "c3_f32_gkeys_s": g("C3_1B_f32_test_shaped", "gkeys_s"), "c3_f32_raw_gkeys_s": g("C3_1B_f32_raw_bits", "gkeys_s"),
        "c3_f64_gkeys_s": g("C3_1B_f64_raw_bits", "gkeys_s"), "c3_i32_gkeys_s": g("C3_1B_i32", "gkeys_s"),
        "c3_i64_gkeys_s": g("C3_1B_i64", "gkeys_s"),
        "c4_zipf_gkeys_s": g("C4_2B_u32_zipf_1.0", "gkeys_s"), "c4_step16_gkeys_s": g("C4_2B_u32_step_16", "gkeys_s"),
        "c4_ascending_gkeys_s": g("C4_2B_u32_ascending", "gkeys_s"), "c4_geometric_gkeys_s": g("C4_2B_u32_geometric_2^-10", "gkeys_s"),
        "records_u32_u32_grecords_s": g("records_4GB_u32_u32", "grecords_s"), "records_u64_u64_grecords_s": g("records_4GB_u64_u64", "grecords_s"),
        "c5_16B_keys_gkeys_s": (c5_line or {}).get("value"), "mgpu_parity": mgpu_parity,
        "cpu_gkeys_s": (cpu or {}).get("value"), "n_errors": len(errors), "failed": failed,
    }
    line["summary"] = {k_: v_ for k_, v_ in summary.items() if v_ is not None or k_ in ("mgpu_parity",)}
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if failed:
        sys.stderr.write(f"bench.py: parity / verification failed in: {failed}\n")
        return 1
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--keys", dest="n", type=int, default=N_KEYS, help="keys per GPU (default: the BASELINE.json workload)")
    ap.add_argument("--cpu-keys", type=int, default=N_KEYS, help="keys of the cpu_baseline leg (default: the whole workload)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-u32", action="store_true", help="skip the secondary u32 measurement")
    ap.add_argument("--no-extra", action="store_true", help="skip the other BASELINE.json configurations (C1, C3, C4, records, C5)")
    ap.add_argument("--no-parity", action="store_true", help="N > 1: skip the untimed multi-GPU parity pre-step")
    ap.add_argument("--no-e2e", action="store_true",
                    help="skip the host-buffer leg (evidence runs at sizes whose pinned host copies would not fit the box)")
    ap.add_argument("--overlap", default="off", choices=["on", "off"],
                    help="N > 1, fused exchange: send the buckets in two pieces and sort the first while the second travels")
    ap.add_argument("--claim", default="off", choices=["on", "off"],
                    help="N > 1, fused exchange: count-free K7 (tiles claim room in the receive buffers through cursors)")
    ap.add_argument("--exchange", default="auto", choices=["auto", "fused", "nccl"],
                    help="N > 1: fused K7 peer stores, NCCL all-to-all, or fused with fallback (default)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
