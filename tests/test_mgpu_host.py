"""CPU, world_size 2 over gloo: the host-side logic of the multi-GPU sort
(radix_sort_b200/mgpu.py) -- sample positions, splitter choice, count exchange,
all-to-all plan -- with a numpy test double standing in for the device steps
(K7 partition and the local sort).  The double lives here, in tests/: the
product's default ops are the CUDA C ABI and refuse to run without a GPU."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from radix_sort_b200 import mgpu  # noqa: E402
from radix_sort_b200.api import Kind  # noqa: E402


class NumpyOps:
    """Test double for mgpu.CudaOps on CPU tensors: rows of 8 bytes = u64 keys."""

    @staticmethod
    def _u64(rows):
        return rows.numpy().reshape(-1).view(np.uint64)

    def gather_samples(self, rows, positions):
        return rows[torch.from_numpy(positions)].contiguous()

    def empty_rows(self, n, key_bytes):
        return torch.empty((n, key_bytes), dtype=torch.uint8)

    def local_sort(self, rows, kind, workspace=None):
        a = self._u64(rows)
        a[...] = np.sort(a, kind="stable")

    def partition(self, rows, out, kind, splitters_host, workspace=None):
        a = self._u64(rows)
        spl = splitters_host.reshape(-1).view(np.uint64)
        dest = np.searchsorted(spl, a, side="right")  # #{j : spl[j] <= key}
        order = np.argsort(dest, kind="stable")
        self._u64(out)[...] = a[order]
        return torch.from_numpy(np.bincount(dest, minlength=len(spl) + 1).astype(np.int64))

    def to_host(self, t):
        return t.numpy()

    # ---- fused path: count, then "peer stores" emulated over gloo.  Every sender
    # ships (offset it computed, bucket) pairs; the receiver writes each bucket at
    # the SENDER's offset, so wrong offsets from mgpu.send_offsets corrupt the result.
    def workspace_for(self, n, kind, workspace=None):
        return None

    def _dest(self, rows, splitters_host):
        a = self._u64(rows)
        spl = splitters_host.reshape(-1).view(np.uint64)
        return a, np.searchsorted(spl, a, side="right")

    def partition_count(self, rows, kind, splitters_host, workspace):
        a, dest = self._dest(rows, splitters_host)
        return torch.from_numpy(np.bincount(dest, minlength=len(splitters_host) + 1).astype(np.int64))

    def partition_exchange(self, rows, kind, splitters_host, workspace, exchange, offsets):
        a, dest = self._dest(rows, splitters_host)
        world = dist.get_world_size()
        mine = [(int(offsets[d]), a[dest == d].copy()) for d in range(world)]
        everyone = [None] * world
        dist.all_gather_object(everyone, mine)
        recv = exchange.recv_rows.numpy().reshape(-1).view(np.uint64)
        me = dist.get_rank()
        for sender in range(world):
            off, part = everyone[sender][me]
            assert off + len(part) <= exchange.capacity
            recv[off: off + len(part)] = part


    # overlapped exchange: one piece = the buckets of `bucket_mask`, bucket j into the
    # receive buffer of rank owners[j]
    def partition_exchange_part(self, rows, kind, splitters_host, workspace, exchange, offsets, owners, bucket_mask,
                                again, channel):
        a, dest = self._dest(rows, splitters_host)
        world = dist.get_world_size()
        mine = [[] for _ in range(world)]
        for j, o in enumerate(owners):
            if (bucket_mask >> j) & 1:
                mine[o].append((int(offsets[j]), a[dest == j].copy()))
        everyone = [None] * world
        dist.all_gather_object(everyone, mine)
        recv = exchange.recv_rows.numpy().reshape(-1).view(np.uint64)
        me = dist.get_rank()
        for sender in range(world):
            for off, part in everyone[sender][me]:
                assert off + len(part) <= exchange.capacity
                recv[off: off + len(part)] = part

    def side_stream(self):
        return None

    def fork(self, side):
        pass

    def join(self, side):
        pass

    def on(self, stream):
        import contextlib
        return contextlib.nullcontext()


class FakeExchange:
    """Stands in for mgpu.PeerExchange: a local receive buffer of `capacity` keys."""

    def __init__(self, capacity, key_bytes=8):
        self.capacity = capacity
        self.recv_rows = torch.full((capacity, key_bytes), 0xEE, dtype=torch.uint8)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, n_per_rank, mode, q, fused=False, capacity=0, overlap=False):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(1000 + rank)
        if mode == "uniform":
            a = rng.integers(0, 2**63, size=n_per_rank, dtype=np.int64)
        elif mode == "few_unique":
            a = rng.integers(0, 5, size=n_per_rank, dtype=np.int64) * 1000
        else:  # ragged: rank 1 holds nothing
            a = rng.integers(0, 2**63, size=n_per_rank if rank == 0 else 0, dtype=np.int64)
        keys = torch.from_numpy(a.copy())
        # int64 tensor holding u64 bit patterns < 2^63 (same order either way)
        exchange = FakeExchange(capacity) if fused else None
        res = mgpu.sort_sharded(keys, Kind.U64, ops=NumpyOps(), samples_per_rank=512, exchange=exchange, overlap=overlap)
        if fused and capacity >= 2 * n_per_rank:  # the result must be a view of the receive buffer
            assert res.keys.data_ptr() == exchange.recv_rows.data_ptr()
        out = res.keys.numpy().copy()
        q.put((rank, a, out, res.send_counts, res.recv_counts))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("mode,fused,capacity,overlap", [
    ("uniform", False, 0, False), ("few_unique", False, 0, False), ("ragged", False, 0, False),
    ("uniform", True, 40_000, False), ("few_unique", True, 40_000, False), ("ragged", True, 40_000, False),
    ("uniform", True, 100, False),  # too small: all ranks fall back together
    # overlapped exchange: ranges cut in two, the halves sent as two pieces
    ("uniform", True, 40_000, True), ("few_unique", True, 40_000, True), ("ragged", True, 40_000, True),
    ("uniform", True, 100, True)])
def test_sort_sharded_world2_gloo(mode, fused, capacity, overlap):
    world, n = 2, 20_000
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, mode, q, fused, capacity, overlap)) for r in range(world)]
    [p.start() for p in procs]
    got = {}
    for _ in range(world):
        r, a, out, sc, rc = q.get(timeout=120)
        got[r] = (a, out, sc, rc)
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    whole_in = np.concatenate([got[r][0] for r in range(world)])
    whole_out = np.concatenate([got[r][1] for r in range(world)])  # shards in rank order
    assert np.array_equal(whole_out, np.sort(whole_in)), "concatenated shards are not the sorted input"
    # the exchange plan is consistent: what r sends to d is what d receives from r
    for r in range(world):
        for d in range(world):
            assert got[r][2][d] == got[d][3][r]
    if mode == "uniform":  # splitters from samples balance the shards
        sizes = [len(got[r][1]) for r in range(world)]
        assert max(sizes) < 1.2 * sum(sizes) / world


def test_sample_positions_and_splitters_are_deterministic():
    p1 = mgpu.sample_positions(1000, 64, seed=7, rank=1)
    p2 = mgpu.sample_positions(1000, 64, seed=7, rank=1)
    assert np.array_equal(p1, p2) and p1.min() >= 0 and p1.max() < 1000
    assert len(mgpu.sample_positions(0, 64, 7, 0)) == 0
    s = np.arange(100)
    assert mgpu.choose_splitters(s, 4) == [25, 50, 75]
    assert mgpu.choose_splitters(s[:0], 4) == []
    send, recv = mgpu.exchange_plan(np.array([[1, 2], [3, 4]]), rank=0)
    assert send == [1, 2] and recv == [1, 3]
    # receive buffers fill in sender order: rank 1's keys start after rank 0's
    cm = np.array([[1, 2, 3], [4, 5, 6], [7, 8, 9]])
    assert mgpu.send_offsets(cm, 0) == [0, 0, 0]
    assert mgpu.send_offsets(cm, 1) == [1, 2, 3]
    assert mgpu.send_offsets(cm, 2) == [5, 7, 9]
    # halved ranges: 2 ranks x 4 sub-ranges; rank d's buffer = [sub-range 2d | sub-range 2d+1], each in sender order
    cm2 = np.array([[1, 2, 3, 4], [5, 6, 7, 8]])
    assert mgpu.halved_plan(cm2, 0) == ([0, 6, 0, 10], [6, 10], [8, 12])
    assert mgpu.halved_plan(cm2, 1) == ([1, 8, 3, 14], [6, 10], [8, 12])


def test_cuda_ops_refuse_to_run_without_a_gpu():
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(Exception) as e:
        mgpu.CudaOps(torch.device("cpu"))
    assert "no CPU fallback" in str(e.value) or "CUDA" in str(e.value)
