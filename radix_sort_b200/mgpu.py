"""Multi-GPU sort: one process per GPU (torch.distributed), keys sharded.

The reference has no multi-device code; its only analogue is the chunk
dimension of the scan (mod.rs:110-120).  LSD passes are global permutations, so
the shards cannot be sorted independently: the keys are MSD-partitioned ONCE by
G-1 splitters, buckets are exchanged all-to-all (NCCL over NVLink / NVSwitch),
and every GPU then LSD-sorts what it received with the single-GPU kernels.
Rank r ends up holding the r-th contiguous range of the global order, so the
concatenation of the shards in rank order is the reference's output.

    1. sample    every rank draws S keys at fixed pseudo-random positions
    2. splitters all_gather(samples) -> every rank sorts the G*S samples (on its
                 GPU, with the library's own sort) and picks the same G-1 quantiles
    3. K7        b200sort_partition_device: stable partition by destination
                 rank + per-destination counts (one onesweep pass keyed by rank)
    4. counts    all_gather(counts) -> receive counts / offsets
    5. exchange  all_to_all_single with split sizes (ncclSend/ncclRecv pairs)
    6. local     b200sort_keys_device on the received keys

Equal keys are bit-identical for key-only sorts, so which GPU holds which copy
of a duplicated key is unobservable; a splitter value that is heavily
duplicated simply lands whole on one rank (load imbalance, not an error).

Fused form (`sort_sharded(..., exchange=PeerExchange(...))`): steps 3-5 become
    3'. b200sort_partition_count_device   one read: per-destination counts
    4'. all_gather(counts) -> for every destination, the offset in ITS receive
        buffer at which this rank's keys start
    5'. b200sort_partition_exchange_device  K7 storing every bucket straight into
        the destination GPU's receive buffer (peer memory over NVLink, mapped by
        torch symmetric memory) + one device-side barrier
so the keys cross NVLink inside the partition kernel's write-out and are never
staged in a send buffer or copied by a separate collective.

Claim form (OPT-IN: `sort_sharded(..., exchange=..., claim=True)`; the counting form above
stays the default): no count phase and no count matrix --
    3*. every rank zeroes its cursor, barrier
    4*. b200sort_partition_claim_device  one read of the shard; every K7 tile claims
        room in the destination's receive buffer with ONE system-scope atomicAdd on a
        cursor that lives on the destination GPU, and stores its keys there (the order
        inside a destination is free: it is sorted next)
    5*. barrier; the cursors now hold every rank's receive count.  One small all_gather
        of (cursor, overflow flags) tells every rank the same thing: if any claim passed
        a receive buffer's capacity nothing of that claim was stored, and ALL ranks take
        the counting form instead (the cursors still hold the true sizes).
A staged variant (second halves claimed into a local staging buffer and moved by the copy
engines under the first local sort) exists in CudaOps.partition_claim(staged=True) and
PeerExchange(staging=True) but nothing drives it yet: it is not part of sort_sharded.

The device steps go through an `ops` object so that the host-side logic (sample
positions, splitter choice, count exchange, offsets) runs under gloo on CPU in
tests/ with a numpy test double; the default ops are the CUDA C ABI and fail
loudly without a GPU.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Optional

import numpy as np

from . import _ffi
from ._ffi import B200SortError, check
from .api import Kind, NUMBER_OF_DIGITS, temp_bytes
from .distr import mix64

MAX_RANKS = 16  # K7 takes at most 15 splitters (include/b200sort.h)


def sample_positions(n_local: int, samples: int, seed: int, rank: int) -> np.ndarray:
    """`samples` pseudo-random positions in [0, n_local), identical on every run."""
    if n_local == 0 or samples == 0:
        return np.zeros(0, dtype=np.int64)
    i = np.arange(samples, dtype=np.uint64)
    u = mix64(np.uint64((seed ^ (0x9E37 * (rank + 1))) & (2**64 - 1)) ^ mix64(i))
    return (u % np.uint64(n_local)).astype(np.int64)


def choose_splitters(sorted_samples, world: int):
    """G-1 quantiles of the globally sorted samples (any indexable sequence)."""
    m = len(sorted_samples)
    if m == 0:
        return []
    return [sorted_samples[min(m - 1, (m * (j + 1)) // world)] for j in range(world - 1)]


def exchange_plan(count_matrix: np.ndarray, rank: int):
    """count_matrix[s][d] = keys rank s sends to rank d.  Returns (send_counts,
    recv_counts) of `rank` as python int lists."""
    cm = np.asarray(count_matrix, dtype=np.int64)
    return [int(x) for x in cm[rank, :]], [int(x) for x in cm[:, rank]]


def send_offsets(count_matrix: np.ndarray, rank: int):
    """Element offset, in every destination's receive buffer, at which the keys
    sent by `rank` start: receive buffers are filled in sender-rank order, so it
    is the number of keys lower ranks send to that destination."""
    cm = np.asarray(count_matrix, dtype=np.int64)
    return [int(x) for x in cm[:rank, :].sum(axis=0)]


def halved_plan(count_matrix: np.ndarray, rank: int):
    """Overlapped exchange: every rank's range is cut in two by one more splitter, so
    count_matrix[s][j] counts the keys rank s holds of SUB-range j (2 * world of them;
    sub-ranges 2d and 2d+1 belong to rank d and lie back to back in d's receive
    buffer, each filled in sender-rank order).  Returns (offsets, first, second):
    offsets[j] = element offset in rank j//2's receive buffer at which `rank`'s
    keys of sub-range j start; first[d] / second[d] = keys rank d receives in its
    first / second sub-range."""
    cm = np.asarray(count_matrix, dtype=np.int64)
    tot = cm.sum(axis=0)
    first, second = tot[0::2], tot[1::2]
    before = cm[:rank, :].sum(axis=0)
    offsets = [int(before[j] + (first[j // 2] if j % 2 else 0)) for j in range(cm.shape[1])]
    return offsets, [int(x) for x in first], [int(x) for x in second]


class PeerExchange:
    """Receive buffers of all ranks of `group`, each `capacity_keys` keys, mapped
    into every process (torch symmetric memory: CUDA VMM allocations exchanged
    between the processes, P2P over NVLink / NVSwitch).  Collective constructor;
    reuse one instance across sorts -- the rendezvous is the expensive part.

    Re-use protocol: a rank's peers start storing into its receive buffer after
    the count all_gather of the NEXT sort, which that rank enqueues on the same
    stream after whatever consumed the previous result; consume the result on
    the stream sort_sharded ran on (or synchronise with it) before the next call."""

    CTRL_BYTES = 512  # per rank, behind the receive area: [0] cursor of the receive buffer, [1..16] cursors of the
                      # staging regions (one per destination), [32] overflow flags (u32) -- int64 words

    def __init__(self, capacity_keys: int, key_bytes: int, device, group=None, staging: bool = False):
        import torch
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        if not torch.cuda.is_available():
            raise B200SortError(3, "peer exchange needs CUDA devices (there is no CPU fallback)")
        self.capacity = int(capacity_keys)
        self.key_bytes = int(key_bytes)
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.recv_bytes = (self.capacity * self.key_bytes + 255) // 256 * 256
        self.buf = symm.empty(self.recv_bytes + self.CTRL_BYTES, dtype=torch.uint8, device=device)
        self.handle = symm.rendezvous(self.buf, self.group)
        self.peer_ptrs = [int(p) for p in self.handle.buffer_ptrs]
        self.recv_rows = self.buf[: self.capacity * self.key_bytes].view(self.capacity, self.key_bytes)
        self.ctrl = self.buf[self.recv_bytes:].view(torch.int64)      # this rank's control block
        self.ctrl.zero_()
        # staging buffer of the second halves (opt-in, not used by sort_sharded: world x capacity / world keys of HBM):
        # one region per destination rank (plain device memory)
        self.stage_cap = self.capacity // max(1, self.world) if staging else 0
        self.stage = (torch.empty(self.world * self.stage_cap * self.key_bytes, dtype=torch.uint8, device=device)
                      if staging and self.stage_cap else None)
        self._peer_views = {}
        torch.cuda.synchronize(device)
        self.handle.barrier(channel=0)  # every rank's cursors are zero before anybody claims

    def cursor_addr(self, rank: int, word: int = 0) -> int:
        """device address (in this process) of control word `word` of `rank`"""
        return self.peer_ptrs[rank] + self.recv_bytes + 8 * word

    def peer_rows(self, rank: int):
        """rank's receive buffer as a (capacity, key_bytes) uint8 tensor of THIS process (peer memory)"""
        import torch
        v = self._peer_views.get(rank)
        if v is None:
            v = self.handle.get_buffer(rank, (self.capacity * self.key_bytes,), torch.uint8, 0).view(self.capacity, self.key_bytes)
            self._peer_views[rank] = v
        return v

    def reset_cursors(self):
        """Claim form: zero this rank's cursors and overflow flags on the current stream, then a barrier -- after
        it every rank's cursors are zero and every rank has reached this point of ITS stream, i.e. has consumed
        the previous result, so peers may start claiming and storing."""
        self.ctrl.zero_()
        self.handle.barrier(channel=0)

    def barrier(self, channel: int = 0):
        """Device-side barrier on the current stream: every rank's stores into
        every receive buffer are complete and visible after it.  Barriers that can
        be in flight at the same time (on different streams) need different channels."""
        self.handle.barrier(channel=channel)


class CudaOps:
    """Device steps on the current CUDA device through the C ABI.  Keys travel as
    (n, key_bytes) uint8 rows: byte tensors index, concatenate and go through
    NCCL whatever the key type (torch has no CUDA indexing for uint64)."""

    def __init__(self, device):
        import torch
        if not torch.cuda.is_available():
            raise B200SortError(3, "multi-GPU sort needs CUDA devices (there is no CPU fallback)")
        self.torch = torch
        self.device = device
        self.lib = _ffi.lib()

    def _stream(self):
        return ctypes.c_void_p(self.torch.cuda.current_stream(self.device).cuda_stream)

    def gather_samples(self, rows, positions: np.ndarray):
        idx = self.torch.from_numpy(positions).to(self.device)
        return rows[idx].contiguous()

    def empty_rows(self, n: int, key_bytes: int):
        return self.torch.empty((n, key_bytes), dtype=self.torch.uint8, device=self.device)

    def local_sort(self, rows, kind: Kind, workspace=None):
        n = rows.shape[0]
        if n == 0:
            return
        need = temp_bytes(n, kind)
        if workspace is None or workspace.numel() < need:
            workspace = self.torch.empty(need, dtype=self.torch.uint8, device=self.device)
        check(self.lib.b200sort_keys_device(rows.data_ptr(), workspace.data_ptr(), workspace.numel(), n, int(kind),
                                            self._stream()))
        workspace.record_stream(self.torch.cuda.current_stream(self.device))

    def partition(self, rows, out, kind: Kind, splitters_host: np.ndarray, workspace=None):
        """K7.  splitters_host: (nsplit, key_bytes) uint8.  Returns the per-destination
        counts as a device int64 tensor."""
        n = rows.shape[0]
        nsplit = int(splitters_host.shape[0])
        counts = self.torch.zeros(nsplit + 1, dtype=self.torch.int64, device=self.device)
        need = temp_bytes(max(n, 1), kind)
        if workspace is None or workspace.numel() < need:
            workspace = self.torch.empty(need, dtype=self.torch.uint8, device=self.device)
        spl = np.ascontiguousarray(splitters_host)
        check(self.lib.b200sort_partition_device(rows.data_ptr(), out.data_ptr(), workspace.data_ptr(), workspace.numel(),
                                                 n, int(kind), spl.ctypes.data if nsplit else None, nsplit,
                                                 counts.data_ptr(), self._stream()))
        workspace.record_stream(self.torch.cuda.current_stream(self.device))
        return counts

    def partition_count(self, rows, kind: Kind, splitters_host: np.ndarray, workspace):
        """K7 count phase; leaves its state in `workspace` for partition_exchange."""
        n = rows.shape[0]
        nsplit = int(splitters_host.shape[0])
        counts = self.torch.zeros(nsplit + 1, dtype=self.torch.int64, device=self.device)
        spl = np.ascontiguousarray(splitters_host)
        check(self.lib.b200sort_partition_count_device(rows.data_ptr(), workspace.data_ptr(), workspace.numel(), n, int(kind),
                                                       spl.ctypes.data if nsplit else None, nsplit, counts.data_ptr(),
                                                       self._stream()))
        return counts

    def partition_exchange(self, rows, kind: Kind, splitters_host: np.ndarray, workspace, exchange, offsets):
        """K7 scatter phase storing into the peers' receive buffers; `offsets`: device
        int64 tensor, element offset of this rank's keys in every destination."""
        n = rows.shape[0]
        nsplit = int(splitters_host.shape[0])
        spl = np.ascontiguousarray(splitters_host)
        bases = (ctypes.c_void_p * (nsplit + 1))(*exchange.peer_ptrs[: nsplit + 1])
        check(self.lib.b200sort_partition_exchange_device(rows.data_ptr(), workspace.data_ptr(), workspace.numel(), n,
                                                          int(kind), spl.ctypes.data if nsplit else None, nsplit, bases,
                                                          offsets.data_ptr(), self._stream()))
        workspace.record_stream(self.torch.cuda.current_stream(self.device))
        exchange.barrier()

    def partition_exchange_part(self, rows, kind: Kind, splitters_host: np.ndarray, workspace, exchange, offsets,
                                owners, bucket_mask: int, again: bool, channel: int):
        """One piece of the K7 scatter phase: only the buckets in `bucket_mask` are stored
        (bucket j into the receive buffer of rank owners[j]), then a barrier on `channel`."""
        n = rows.shape[0]
        nsplit = int(splitters_host.shape[0])
        spl = np.ascontiguousarray(splitters_host)
        bases = (ctypes.c_void_p * (nsplit + 1))(*[exchange.peer_ptrs[o] for o in owners])
        check(self.lib.b200sort_partition_exchange_part_device(rows.data_ptr(), workspace.data_ptr(), workspace.numel(), n,
                                                               int(kind), spl.ctypes.data if nsplit else None, nsplit, bases,
                                                               offsets.data_ptr(), int(bucket_mask), 1 if again else 0,
                                                               self._stream()))
        workspace.record_stream(self.torch.cuda.current_stream(self.device))
        exchange.barrier(channel)

    def partition_claim(self, rows, kind: Kind, splitters_host: np.ndarray, workspace, exchange, staged: bool):
        """K7 in claim mode (b200sort_partition_claim_device).  staged=False: bucket r is rank r's range, stored
        into rank r's receive buffer.  staged=True: 2 * world buckets, bucket 2 d -> rank d's receive buffer,
        bucket 2 d + 1 -> region d of the local staging buffer."""
        n = rows.shape[0]
        nsplit = int(splitters_host.shape[0])
        spl = np.ascontiguousarray(splitters_host)
        nb = nsplit + 1
        me = exchange.handle.rank
        kb = exchange.key_bytes
        if staged:
            stage0 = exchange.stage.data_ptr()
            bases = [exchange.peer_ptrs[j // 2] if j % 2 == 0 else stage0 + (j // 2) * exchange.stage_cap * kb for j in range(nb)]
            cursors = [exchange.cursor_addr(j // 2, 0) if j % 2 == 0 else exchange.cursor_addr(me, 1 + j // 2) for j in range(nb)]
            caps = [exchange.capacity if j % 2 == 0 else exchange.stage_cap for j in range(nb)]
        else:
            bases = [exchange.peer_ptrs[j] for j in range(nb)]
            cursors = [exchange.cursor_addr(j, 0) for j in range(nb)]
            caps = [exchange.capacity] * nb
        c_bases = (ctypes.c_void_p * nb)(*bases)
        c_cur = (ctypes.c_uint64 * nb)(*cursors)
        c_cap = (ctypes.c_uint64 * nb)(*caps)
        flags = ctypes.c_void_p(exchange.cursor_addr(me, 32))
        check(self.lib.b200sort_partition_claim_device(rows.data_ptr(), workspace.data_ptr(), workspace.numel(), n, int(kind),
                                                       spl.ctypes.data if nsplit else None, nsplit, c_bases, c_cur, c_cap, flags,
                                                       self._stream()))
        workspace.record_stream(self.torch.cuda.current_stream(self.device))

    def device_splitters(self, rows, kind: Kind, samples: int, seed: int, rank: int, world: int, parts: int, group):
        """`parts - 1` quantiles of a global sample, picked on the device: ONE all_gather of a fixed-size block
        (samples + their count), one sort, one tiny copy to the host.  Ranks with fewer than `samples` keys pad
        with the largest key of the kind's order; the pads sort behind every real key and the quantile indices
        are taken over the real ones only.  Returns ((parts - 1, key_bytes) uint8 host array, total samples)."""
        import torch.distributed as dist
        torch = self.torch
        kb = rows.shape[1]
        n_local = int(rows.shape[0])
        key = (n_local, samples, seed, rank)
        cache = getattr(self, "_pos_cache", None)
        if cache is None or cache[0] != key:
            pos = sample_positions(n_local, min(samples, n_local) if n_local else 0, seed, rank)
            idx = torch.from_numpy(pos).to(self.device)
            self._pos_cache = cache = (key, idx)
        idx = cache[1]
        m = int(idx.shape[0])
        # block = samples rows (pads behind the real ones) + 8 bytes holding the number of real rows
        flat = torch.empty(samples * kb + 8, dtype=torch.uint8, device=self.device)
        body = flat[: samples * kb].view(samples, kb)
        if m < samples:
            pad = np.full(kb, 0xFF, dtype=np.uint8)
            if kind in (Kind.I8, Kind.I16, Kind.I32, Kind.I64, Kind.I128, Kind.F32, Kind.F64):
                pad[kb - 1] = 0x7F  # little-endian: the largest signed value / the positive NaN with the largest payload
            body[m:] = torch.from_numpy(pad).to(self.device)
        if m:
            body[:m] = rows[idx]
        flat[samples * kb:] = torch.tensor([m], dtype=torch.int64).view(torch.uint8).to(self.device)
        allb = torch.empty((world, samples * kb + 8), dtype=torch.uint8, device=self.device)
        dist.all_gather_into_tensor(allb, flat, group=group)
        total = allb[:, samples * kb:].contiguous().view(torch.int64).sum()
        pool = allb[:, : samples * kb].reshape(world * samples, kb).contiguous()
        self.local_sort(pool, kind)
        j = torch.arange(1, parts, dtype=torch.int64, device=self.device)
        pick_idx = torch.clamp((total * j) // parts, max=max(world * samples - 1, 0))
        picks = pool[pick_idx]
        packed = torch.cat([picks.reshape(-1), total.reshape(1).view(torch.uint8)]).cpu().numpy()
        tot = int(packed[-8:].view(np.int64)[0])
        return np.ascontiguousarray(packed[:-8].reshape(parts - 1, kb)), tot

    def side_stream(self):
        """A second stream (created once) for the piece of the exchange that overlaps the first local sort."""
        if getattr(self, "_side", None) is None:
            self._side = self.torch.cuda.Stream(self.device)
        return self._side

    def fork(self, side):
        """`side` waits for everything enqueued so far on the current stream."""
        side.wait_stream(self.torch.cuda.current_stream(self.device))

    def on(self, stream):
        return self.torch.cuda.stream(stream)

    def join(self, side):
        """The current stream waits for everything enqueued so far on `side`."""
        self.torch.cuda.current_stream(self.device).wait_stream(side)

    def workspace_for(self, n: int, kind: Kind, workspace=None):
        need = temp_bytes(max(n, 1), kind)
        if workspace is None or workspace.numel() < need:
            workspace = self.torch.empty(need, dtype=self.torch.uint8, device=self.device)
        return workspace

    def to_host(self, t) -> np.ndarray:
        return t.cpu().numpy()

    def mark(self):
        ev = self.torch.cuda.Event(enable_timing=True)
        ev.record(self.torch.cuda.current_stream(self.device))
        return ev

    def elapsed_ms(self, a, b) -> float:
        b.synchronize()
        return float(a.elapsed_time(b))


@dataclass
class ShardedResult:
    keys: object          # this rank's range of the global order (tensor, length n_out)
    n_out: int
    send_counts: list
    recv_counts: list
    splitters: np.ndarray  # raw key bit patterns, host


def sort_sharded(keys, kind, *, group=None, ops=None, samples_per_rank: int = 1 << 14, seed: int = 0xB200,
                 workspace=None, send_buffer=None, exchange=None, timings: Optional[dict] = None,
                 overlap: bool = False, workspace2=None, claim: bool = False) -> ShardedResult:
    """Sort keys sharded over the ranks of `group`.  `keys`: this rank's shard
    (contiguous CUDA tensor; (n, 2) uint64 for 128-bit kinds).  Collective: every
    rank of the group must call it.  Returns this rank's slice of the sorted order
    (same dtype as `keys`).

    exchange: a PeerExchange -> fused path (K7 stores into the peers' receive
    buffers; the result is a view of this rank's receive buffer).  A sort whose
    largest shard would not fit the exchange's capacity falls back, on all ranks
    together, to the all-to-all path.  timings: dict filled with per-phase
    milliseconds (device time on the current stream).

    overlap (fused path, world <= 8): every rank's range is cut in two by one more
    splitter; the first halves are exchanged, then the second halves cross NVLink on
    a side stream WHILE every rank already sorts its first half (workspace2: a second
    sort workspace, the partition state in `workspace` is still in use then).

    claim (fused path, without overlap): the count-free exchange of the module docstring -- K7 tiles claim
    room in the destinations' receive buffers through cursors instead of counting first.  The result has no
    per-sender counts then: send_counts is None, recv_counts[d] = keys rank d received.  Falls back to the
    counting form, on all ranks together, when a receive buffer would overflow."""
    import torch
    import torch.distributed as dist
    k = Kind(kind)
    kb = NUMBER_OF_DIGITS[k]
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    if world > MAX_RANKS:
        raise B200SortError(1, f"at most {MAX_RANKS} ranks (K7 takes 15 splitters)")
    if ops is None:
        ops = CudaOps(keys.device)
    tail = tuple(keys.shape[1:])
    # one row of kb bytes per key
    rows = keys.view(torch.uint8).reshape(-1, kb) if keys.numel() else torch.empty((0, kb), dtype=torch.uint8, device=keys.device)
    n_local = int(rows.shape[0])

    def typed(r):
        return r.reshape(-1).view(keys.dtype).reshape((-1,) + tail)

    marks = []

    def mark(name):
        if timings is not None and hasattr(ops, "mark"):
            marks.append((name, ops.mark()))

    def finish():
        if timings is not None and marks:
            for (_, a), (name, b) in zip(marks[:-1], marks[1:]):
                timings[name] = timings.get(name, 0.0) + ops.elapsed_ms(a, b)

    mark("start")
    if world == 1:
        ops.local_sort(rows, k, workspace)
        return ShardedResult(keys, n_local, [n_local], [n_local], np.zeros((0, kb), dtype=np.uint8))

    # 1-2. splitters from a global sample
    pos = sample_positions(n_local, samples_per_rank, seed, rank)
    mine = ops.gather_samples(rows, pos)
    cnt = torch.tensor([mine.shape[0]], dtype=torch.int64, device=rows.device)
    cnts = [torch.zeros_like(cnt) for _ in range(world)]
    dist.all_gather(cnts, cnt, group=group)
    sizes = [int(c.item()) for c in cnts]
    padded = ops.empty_rows(samples_per_rank, kb)
    padded[: mine.shape[0]] = mine
    gathered = [torch.empty_like(padded) for _ in range(world)]
    dist.all_gather(gathered, padded, group=group)
    allsamp = torch.cat([g[:s] for g, s in zip(gathered, sizes)], dim=0).contiguous()
    ops.local_sort(allsamp, k)
    samp_host = ops.to_host(allsamp)
    halves = exchange is not None and overlap and 2 * world <= MAX_RANKS and hasattr(ops, "partition_exchange_part")
    picks = choose_splitters(samp_host, 2 * world if halves else world)
    if len(picks) != (2 * world if halves else world) - 1:  # no keys anywhere
        return ShardedResult(keys[:0], 0, [0] * world, [0] * world, np.zeros((0, kb), dtype=np.uint8))
    fine = np.ascontiguousarray(np.stack(picks))
    splitters = np.ascontiguousarray(fine[1::2]) if halves else fine  # the world - 1 range boundaries
    mark("splitters")

    if exchange is not None and claim and not halves and hasattr(ops, "partition_claim"):
        ws = ops.workspace_for(n_local, k, workspace)
        exchange.reset_cursors()
        ops.partition_claim(rows, k, splitters, ws, exchange, staged=False)
        exchange.barrier()
        mine2 = exchange.ctrl[[0, 32]].clone()  # [keys claimed in my receive buffer, buckets I could not store (bits)]
        both = [torch.zeros_like(mine2) for _ in range(world)]
        dist.all_gather(both, mine2, group=group)
        info = np.stack([ops.to_host(b) for b in both])  # world x 2, the same on every rank
        mark("exchange")
        if not info[:, 1].any() and int(info[:, 0].max()) <= exchange.capacity:
            n_out = int(info[rank, 0])
            recv = exchange.recv_rows[:n_out]
            ops.local_sort(recv, k, ws)
            mark("local_sort")
            finish()
            return ShardedResult(typed(recv), n_out, None, [int(x) for x in info[:, 0]], splitters)
        # a receive buffer would overflow: the counting form below decides what to do, identically on every rank

    if halves:
        # 3''. count per SUB-range  4''. count matrix -> offsets of my keys in every sub-range
        ws = ops.workspace_for(n_local, k, workspace)
        counts = ops.partition_count(rows, k, fine, ws)
        rws = [torch.zeros_like(counts) for _ in range(world)]
        dist.all_gather(rws, counts, group=group)
        cm2 = np.stack([ops.to_host(r) for r in rws])            # world x 2*world
        cm = cm2[:, 0::2] + cm2[:, 1::2]                          # world x world, per destination rank
        send_counts, recv_counts = exchange_plan(cm, rank)
        n_out = sum(recv_counts)
        mark("count")
        if int(cm.sum(axis=0).max()) <= exchange.capacity:  # same decision on every rank
            offs, first, second = halved_plan(cm2, rank)
            offsets = torch.tensor(offs, dtype=torch.int64).to(rows.device)
            owners = [j // 2 for j in range(2 * world)]
            even = sum(1 << j for j in range(0, 2 * world, 2))
            # 5''. first halves: K7 + barrier on this stream
            ops.partition_exchange_part(rows, k, fine, ws, exchange, offsets, owners, even, False, 0)
            mark("exchange")
            # second halves on the side stream (after the first K7: they share its state in ws) ...
            side = ops.side_stream()
            ops.fork(side)
            with ops.on(side):
                ops.partition_exchange_part(rows, k, fine, ws, exchange, offsets, owners, even << 1, True, 1)
            # ... while this stream sorts the first half, then the second once it is complete
            n_a, n_b = first[rank], second[rank]
            ws2 = ops.workspace_for(max(n_a, n_b), k, workspace2)
            ops.local_sort(exchange.recv_rows[:n_a], k, ws2)
            ops.join(side)
            ops.local_sort(exchange.recv_rows[n_a:n_a + n_b], k, ws2)
            mark("local_sort")
            finish()
            return ShardedResult(typed(exchange.recv_rows[:n_out]), n_out, send_counts, recv_counts, splitters)
        # would overflow the receive buffers: the unfused path below, with the range splitters only

    elif exchange is not None:
        # 3'. count  4'. count matrix -> where my keys start in every receive buffer
        ws = ops.workspace_for(n_local, k, workspace)
        counts = ops.partition_count(rows, k, splitters, ws)
        rws = [torch.zeros_like(counts) for _ in range(world)]
        dist.all_gather(rws, counts, group=group)
        cm = np.stack([ops.to_host(r) for r in rws])
        send_counts, recv_counts = exchange_plan(cm, rank)
        n_out = sum(recv_counts)
        mark("count")
        if int(cm.sum(axis=0).max()) <= exchange.capacity:  # same decision on every rank
            offsets = torch.tensor(send_offsets(cm, rank), dtype=torch.int64).to(rows.device)
            # 5'. K7 scatter into the peers' receive buffers + barrier
            ops.partition_exchange(rows, k, splitters, ws, exchange, offsets)
            mark("exchange")
            recv = exchange.recv_rows[:n_out]
            ops.local_sort(recv, k, ws)
            mark("local_sort")
            finish()
            return ShardedResult(typed(recv), n_out, send_counts, recv_counts, splitters)

    # 3. K7: partition by destination rank
    if send_buffer is None or send_buffer.numel() * send_buffer.element_size() < n_local * kb:
        send = ops.empty_rows(n_local, kb)
    else:
        send = send_buffer.view(torch.uint8).reshape(-1)[: n_local * kb].reshape(n_local, kb)
    counts = ops.partition(rows, send, k, splitters, workspace)
    mark("partition")

    # 4. count matrix
    rws = [torch.zeros_like(counts) for _ in range(world)]
    dist.all_gather(rws, counts, group=group)
    cm = np.stack([ops.to_host(r) for r in rws])
    send_counts, recv_counts = exchange_plan(cm, rank)
    n_out = sum(recv_counts)

    # 5. all-to-all: split sizes count rows (= keys)
    recv = ops.empty_rows(n_out, kb)
    dist.all_to_all_single(recv, send, output_split_sizes=recv_counts, input_split_sizes=send_counts, group=group)
    mark("all_to_all")

    # 6. local LSD sort of the received range
    ops.local_sort(recv, k, workspace)
    mark("local_sort")
    finish()
    return ShardedResult(typed(recv), n_out, send_counts, recv_counts, splitters)


class _DeviceSpan:
    """`nbytes` bytes of device memory at `ptr`, for torch.as_tensor (CUDA array interface)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3}


class Communicator:
    """All GPUs of ONE process behind one call: `b200sort_mgpu_create` /
    `_keys_device` / `_destroy` (include/b200sort.h).  The reference's caller is a
    single process (`slice.radix_sort()`, mod.rs:18-20,61), so this is the form a
    shim for it would use when the shards already sit on several GPUs: no launcher,
    no NCCL -- the exchange is K7 storing into peer memory, the counts travel
    through the host.

        comm = Communicator([0, 1, 2, 3])
        outs = comm.sort([shard_on_gpu0, ..., shard_on_gpu3], Kind.U64)

    outs[r] is the r-th contiguous range of the global order, on device r."""

    def __init__(self, devices=None):
        import torch
        if devices is None:
            devices = list(range(torch.cuda.device_count()))
        self.devices = [int(d) for d in devices]
        for d in sorted(set(self.devices)):
            if 0 <= d < torch.cuda.device_count():  # (the library reports the others)
                torch.zeros(1, device=f"cuda:{d}")  # make sure every device has its primary context
        arr = (ctypes.c_int * len(self.devices))(*self.devices)
        self._h = ctypes.c_void_p()
        check(_ffi.lib().b200sort_mgpu_create(ctypes.byref(self._h), len(self.devices), arr))

    def sort(self, shards, kind, copy: bool = True):
        """shards[r]: contiguous CUDA tensor on devices[r] ((n, 2) uint64 for 128-bit
        kinds).  Returns one tensor per device.  copy=False returns views of the
        communicator's own buffers, valid until its next sort."""
        import torch
        k = Kind(kind)
        kb = NUMBER_OF_DIGITS[k]
        G = len(self.devices)
        if len(shards) != G:
            raise B200SortError(1, f"{G} devices but {len(shards)} shards")
        for t, d in zip(shards, self.devices):
            if not (t.is_cuda and t.device.index == d and t.is_contiguous()):
                raise B200SortError(1, "every shard must be a contiguous tensor on its communicator device")
            if t.numel() * t.element_size() % kb:
                raise B200SortError(1, "shard size is not a whole number of keys")
        for d in sorted(set(self.devices)):  # the library reads every shard on its own streams, on every device
            torch.cuda.synchronize(d)
        ptrs = (ctypes.c_void_p * G)(*[t.data_ptr() if t.numel() else None for t in shards])
        n_in = (ctypes.c_size_t * G)(*[t.numel() * t.element_size() // kb for t in shards])
        outs = (ctypes.c_void_p * G)()
        n_out = (ctypes.c_size_t * G)()
        check(_ffi.lib().b200sort_mgpu_keys_device(self._h, ptrs, n_in, int(k), outs, n_out))
        res = []
        for r, (t, d) in enumerate(zip(shards, self.devices)):
            tail = tuple(t.shape[1:])
            if n_out[r] == 0:
                res.append(torch.empty((0,) + tail, dtype=t.dtype, device=t.device))
                continue
            raw = torch.as_tensor(_DeviceSpan(int(outs[r]), int(n_out[r]) * kb), device=f"cuda:{d}")
            view = raw.view(t.dtype).reshape((-1,) + tail)
            res.append(view.clone() if copy else view)
        return res

    def sort_host(self, keys, kind) -> None:
        """`keys`: a C-contiguous writable numpy array in HOST memory, sorted in place by
        all devices of the communicator (`b200sort_mgpu_keys_host`): the drop-in call
        for `slice.radix_sort()` on a box with several GPUs."""
        k = Kind(kind)
        if not (isinstance(keys, np.ndarray) and keys.flags["C_CONTIGUOUS"] and keys.flags["WRITEABLE"]):
            raise B200SortError(1, "keys must be a C-contiguous writable numpy array")
        if keys.nbytes % NUMBER_OF_DIGITS[k]:
            raise B200SortError(1, "array size is not a whole number of keys")
        check(_ffi.lib().b200sort_mgpu_keys_host(self._h, keys.ctypes.data, keys.nbytes // NUMBER_OF_DIGITS[k], int(k)))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            _ffi.lib().b200sort_mgpu_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001 -- interpreter shutdown
            pass
