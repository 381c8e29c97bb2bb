"""Worker of tests/test_mgpu_gpu.py: one rank of a sharded sort under torchrun.
Checks the fused peer-store exchange, its two-piece (overlapped) form and the
NCCL all-to-all path against numpy.  bench.py calls run_cases() /
communicator_case() as its untimed parity pre-step on N >= 2 GPUs, so that the
multi-GPU paths are checked wherever the driver runs the scaling bench."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import radix_sort_b200 as rs  # noqa: E402
from radix_sort_b200 import mgpu  # noqa: E402


def run_cases(dist, dev, rank, world):
    """every rank: sharded sorts of small inputs, gathered and compared with numpy's sort"""
    exchange = mgpu.PeerExchange(400_000, 8, dev)
    for case, n in (("uniform", 250_000), ("few_unique", 250_000), ("ragged", 250_000 if rank == 0 else 0),
                    ("overflow", 500_000)):  # overflow: shards exceed the exchange capacity -> fallback path
        rng = np.random.default_rng(77 + rank)
        a = rng.integers(0, 2**64, size=n, dtype=np.uint64)
        if case == "few_unique":
            a = (a % 5) * 1000
        # (fused exchange, overlapped in two pieces, count-free claim form); the last is the NCCL all-to-all path
        for use_exchange, overlap, claim in ((True, False, False), (True, True, False), (True, False, True), (False, False, False)):
            if overlap and 2 * world > mgpu.MAX_RANKS:
                continue
            keys = torch.from_numpy(a.copy()).to(dev)
            res = mgpu.sort_sharded(keys, rs.Kind.U64, exchange=exchange if use_exchange else None, overlap=overlap, claim=claim)
            # (few_unique: five distinct values, so one rank's range can legitimately exceed the capacity on >= 4 ranks
            # and the claim form hands over to the counting form, like "overflow" always does)
            if claim and case in ("uniform", "ragged"):
                assert res.send_counts is None, "the claim form must have run (it has no per-sender counts)"
            mine = res.keys.cpu().numpy().copy()
            gathered = [None] * world
            dist.all_gather_object(gathered, (a, mine))
            whole_in = np.concatenate([g[0] for g in gathered])
            whole_out = np.concatenate([g[1] for g in gathered])
            assert np.array_equal(whole_out, np.sort(whole_in)), (case, use_exchange, overlap, claim)
            torch.cuda.synchronize()
            dist.barrier()
    del exchange


def communicator_case(devices):
    """ONE process over `devices` (b200sort_mgpu_*): device shards and one host slice, against numpy"""
    comm = mgpu.Communicator(devices)
    try:
        g = len(devices)
        for dt, kind in ((np.uint64, rs.Kind.U64), (np.float32, rs.Kind.F32), (np.int32, rs.Kind.I32)):
            rng = np.random.default_rng(91)
            parts = []
            for r in range(g):
                raw = rng.integers(0, 256, size=(300_000 + 4099 * r) * np.dtype(dt).itemsize, dtype=np.uint8)
                parts.append(raw.view(dt).copy())
            outs = comm.sort([torch.from_numpy(p).to(f"cuda:{d}") for p, d in zip(parts, devices)], kind)
            got = np.concatenate([o.cpu().numpy() for o in outs])
            whole = np.concatenate(parts)
            u = np.dtype(f"u{np.dtype(dt).itemsize}")
            if kind == rs.Kind.F32:   # total_cmp order (radix_digits.rs:103-113): compare bit patterns under the transform
                bits = whole.view(np.uint32)
                keyed = np.where(bits >> 31, ~bits, bits | np.uint32(0x80000000))
                want = whole[np.argsort(keyed, kind="stable")]
            else:
                want = np.sort(whole)
            assert got.view(u).tobytes() == want.view(u).tobytes(), f"communicator: {kind.name} differs"
        # one pinned host slice over all devices (b200sort_mgpu_keys_host)
        _lib = rs._ffi.lib()
        _lib.b200sort_mgpu_set_host_min_bytes(0)
        try:
            rng = np.random.default_rng(92)
            a = rng.integers(0, 2**64, size=1_000_003, dtype=np.uint64)
            buf = torch.empty(a.size, dtype=torch.uint64, pin_memory=True)
            view = buf.numpy()
            view[...] = a
            comm.sort_host(view, rs.Kind.U64)
            assert np.array_equal(view, np.sort(a)), "communicator: host slice differs"
        finally:
            _lib.b200sort_mgpu_set_host_min_bytes(32 << 20)
    finally:
        comm.close()


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        run_cases(dist, dev, rank, world)
        if rank == 0:
            print("MGPU_OK", flush=True)
    finally:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
