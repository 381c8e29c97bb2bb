"""b200sort_keys_device_noalloc: the device sort for callers that cannot supply a workspace.

Some users of the library have no way to allocate device memory: the Rust shim binds
libb200sort.so and nothing else (there is no cudaMalloc in its world, it only receives
device pointers), and a plugin inside a process whose own allocator has reserved all of
HBM must not call cudaMalloc behind the owner's back.  For them the library carries a
16 MiB scratch array in its own image (B200SORT_NOALLOC_SCRATCH_BYTES) and this entry point
sorts with it.  The array is shared between calls and never cleared in between, so the
interesting case is the SECOND call: it runs on whatever the first one left behind.

The checks go through the C ABI and compare, bit for bit (integer / byte work: no
tolerance), with b200sort_keys_device given a fresh caller-owned workspace on the same
input.  The file is named so that pytest collects it last: a failure here must not hide
the rest of the suite under `-x`.
"""
import ctypes

import numpy as np
import pytest

import radix_sort_b200 as rs
from radix_sort_b200 import _ffi

pytestmark = pytest.mark.gpu

SCRATCH_BYTES = 16 << 20  # B200SORT_NOALLOC_SCRATCH_BYTES, include/b200sort.h


def _keys(kind, n, seed):
    rng = np.random.default_rng(seed)
    if kind in ("f32", "f64"):
        bits = rng.integers(0, 2 ** 63, size=n, dtype=np.uint64)
        a = (bits >> (0 if kind == "f64" else 32)).astype(np.uint64 if kind == "f64" else np.uint32)
        return a.view(np.float64 if kind == "f64" else np.float32)  # every bit pattern: NaNs, -0.0, subnormals
    dt = {"u8": np.uint8, "i16": np.int16, "u32": np.uint32, "i32": np.int32, "u64": np.uint64, "i64": np.int64}[kind]
    info = np.iinfo(dt)
    return rng.integers(info.min, info.max, size=n, dtype=dt, endpoint=True)


def _bits(t):
    import torch
    return t.view(torch.uint8).cpu().numpy().tobytes()


def _sort_with_own_workspace(lib, a, kind, cuda):
    """the existing entry point, workspace allocated by the caller (torch here)"""
    import torch
    t = torch.from_numpy(a.copy()).to(cuda)
    ws = torch.empty(lib.b200sort_temp_bytes(a.size, int(kind)) + 256, dtype=torch.uint8, device=cuda)
    ws_ptr = (ws.data_ptr() + 255) // 256 * 256
    _ffi.check(lib.b200sort_keys_device(t.data_ptr(), ws_ptr, ws.numel() - 256, a.size, int(kind),
                                        torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    return t


def _sort_noalloc(lib, a, kind, cuda):
    """the new entry point: keys in, keys out, nothing else"""
    import torch
    t = torch.from_numpy(a.copy()).to(cuda)
    _ffi.check(lib.b200sort_keys_device_noalloc(t.data_ptr(), a.size, int(kind), torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    return t


# (kind, n): sizes above 4096 so that the K1/K2/K3 (or, for 1- and 2-byte keys, the counting)
# path runs and the scratch is really used; the first is close to the limit for 8-byte keys,
# the later ones are smaller and of other widths, so they find foreign leftovers in the array
CALLS = [("u64", 1_900_000), ("u32", 300_001), ("f64", 70_000), ("i16", 1_000_003), ("u8", 5_000_000),
         ("i32", 3_900_000), ("f32", 4097), ("i64", 123_457)]


def test_two_calls_in_a_row_match_the_caller_workspace_entry_point(cuda):
    import torch
    lib = _ffi.lib()
    with torch.cuda.device(cuda):
        for round_ in range(2):  # the whole sequence twice: every shape also runs on its own leftovers
            for i, (kind, n) in enumerate(CALLS):
                k = rs.Kind[kind.upper()]
                assert lib.b200sort_temp_bytes(n, int(k)) <= SCRATCH_BYTES
                a = _keys(kind, n, seed=1000 * round_ + i)
                want = _sort_with_own_workspace(lib, a, k, cuda)
                got = _sort_noalloc(lib, a, k, cuda)
                assert _bits(got) == _bits(want), f"call {i} of round {round_} ({kind}, n={n}) differs"
                # and straight again on the same input, now on this very sort's leftovers
                again = _sort_noalloc(lib, a, k, cuda)
                assert _bits(again) == _bits(want), f"repeat of call {i} of round {round_} ({kind}, n={n}) differs"


def test_back_to_back_without_a_sync_in_between(cuda):
    """two sorts enqueued on one stream with no host synchronisation between them: stream order
    is all the protection the shared array needs"""
    import torch
    lib = _ffi.lib()
    with torch.cuda.device(cuda):
        a = _keys("u64", 1_500_000, seed=7)
        b = _keys("u32", 2_000_000, seed=8)
        want_a = _sort_with_own_workspace(lib, a, rs.Kind.U64, cuda)
        want_b = _sort_with_own_workspace(lib, b, rs.Kind.U32, cuda)
        ta, tb = torch.from_numpy(a.copy()).to(cuda), torch.from_numpy(b.copy()).to(cuda)
        st = torch.cuda.current_stream().cuda_stream
        _ffi.check(lib.b200sort_keys_device_noalloc(ta.data_ptr(), a.size, int(rs.Kind.U64), st))
        _ffi.check(lib.b200sort_keys_device_noalloc(tb.data_ptr(), b.size, int(rs.Kind.U32), st))
        torch.cuda.synchronize()
        assert _bits(ta) == _bits(want_a)
        assert _bits(tb) == _bits(want_b)


def test_sizes_beyond_the_scratch_are_refused_and_nothing_is_touched(cuda):
    import torch
    lib = _ffi.lib()
    with torch.cuda.device(cuda):
        n = 3_000_000  # 8-byte keys: 24 MB of ping-pong buffer alone
        assert lib.b200sort_temp_bytes(n, int(rs.Kind.U64)) > SCRATCH_BYTES
        a = _keys("u64", n, seed=3)
        t = torch.from_numpy(a.copy()).to(cuda)
        before = lib.b200sort_launch_count()
        rc = lib.b200sort_keys_device_noalloc(t.data_ptr(), n, int(rs.Kind.U64), None)
        assert rc == 1  # B200SORT_ERR_INVALID_ARG
        msg = lib.b200sort_last_error().decode()
        assert "B200SORT_NOALLOC_SCRATCH_BYTES" in msg and str(SCRATCH_BYTES) in msg
        assert lib.b200sort_launch_count() == before
        torch.cuda.synchronize()
        assert _bits(t) == a.tobytes()
        # absurd sizes are refused by arithmetic, before any layout is computed
        assert lib.b200sort_keys_device_noalloc(t.data_ptr(), ctypes.c_size_t(2 ** 62).value, int(rs.Kind.U64), None) == 1
        # the largest n that fits is accepted
        lo, hi = 4096, n
        while lo < hi:
            m = (lo + hi + 1) // 2
            if lib.b200sort_temp_bytes(m, int(rs.Kind.U64)) <= SCRATCH_BYTES:
                lo = m
            else:
                hi = m - 1
        b = a[:lo]
        want = _sort_with_own_workspace(lib, b, rs.Kind.U64, cuda)
        got = _sort_noalloc(lib, b, rs.Kind.U64, cuda)
        assert _bits(got) == _bits(want)
        # empty input: a no-op, like the other entry points
        assert lib.b200sort_keys_device_noalloc(None, 0, int(rs.Kind.U64), None) == 0


def test_the_call_allocates_no_device_memory(cuda):
    """free HBM as the driver reports it is the same before and after a run of calls (the first
    call, which resolves the array's address, is made before the measurement)"""
    import torch
    lib = _ffi.lib()
    with torch.cuda.device(cuda):
        a = _keys("u64", 1_000_000, seed=5)
        t = torch.from_numpy(a).to(cuda)
        st = torch.cuda.current_stream().cuda_stream
        _ffi.check(lib.b200sort_keys_device_noalloc(t.data_ptr(), a.size, int(rs.Kind.U64), st))
        torch.cuda.synchronize()
        free0, _ = torch.cuda.mem_get_info()
        for _ in range(8):
            _ffi.check(lib.b200sort_keys_device_noalloc(t.data_ptr(), a.size, int(rs.Kind.U64), st))
        torch.cuda.synchronize()
        free1, _ = torch.cuda.mem_get_info()
        assert free1 == free0
