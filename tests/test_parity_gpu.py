"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle and the
independent numpy statement of the reference's test oracle, bit for bit.

The first block restates the reference's 15 tests (src/radix_sort/tests.rs) one
for one -- same sizes (1e6), same input families, same equality -- with fixed
seeds instead of thread_rng().  The rest covers what the reference leaves
untested (tests.rs has no n==0, tiny n, duplicates, presorted, negative NaN...).
"""
import ctypes
import threading

import numpy as np
import pytest

import radix_sort_b200 as rs
from oracle import cpu_oracle
from radix_sort_b200 import _ffi, distr, verify
from tests import _ref

pytestmark = pytest.mark.gpu

N_REF = 1_000_000  # 1e6 as usize, tests.rs:27,36,45,...


def _kind(name):
    return rs.Kind[name.upper()]


def _check_host(a, kind):
    """sort a copy through b200sort_keys_host; compare with oracle and numpy."""
    want_np = _ref.expected_sorted(a, kind)
    want_or = a.copy()
    cpu_oracle.sort0(want_or, kind)
    assert _ref.same_bits(want_or, want_np), "oracle and numpy disagree"
    got = a.copy()
    rs.radix_sort(got, _kind(kind))
    assert _ref.same_bits(got, want_or), f"{kind}: GPU result differs from the oracle"


def _to_device(a, cuda):
    import torch
    return torch.from_numpy(a).to(cuda)


def _check_device(a, kind, cuda, **kw):
    import torch
    want = a.copy()
    cpu_oracle.sort0(want, kind)
    t = _to_device(a, cuda)
    rs.radix_sort(t, _kind(kind), **kw)
    torch.cuda.synchronize()
    got = t.cpu().numpy()
    assert _ref.same_bits(got, want), f"{kind}: GPU device-path result differs from the oracle"


# ---- the reference's 15 tests, restated (tests.rs:25-187) -------------------------

@pytest.mark.parametrize("kind", ["u8", "u16", "u32", "u64", "u128", "u64"],
                         ids=["radix_sort_u8", "radix_sort_u16", "radix_sort_u32", "radix_sort_u64",
                              "radix_sort_u128", "radix_sort_usize"])
def test_ref_unsigned(kind, cuda):  # tests.rs:25-77
    _check_host(_ref.random_keys(kind, N_REF, seed=0xB200 + len(kind)), kind)


@pytest.mark.parametrize("kind", ["i8", "i16", "i32", "i64", "i128", "i64"],
                         ids=["radix_sort_i8", "radix_sort_i16", "radix_sort_i32", "radix_sort_i64",
                              "radix_sort_i128", "radix_sort_isize"])
def test_ref_signed(kind, cuda):  # tests.rs:79-131
    _check_host(_ref.random_keys(kind, N_REF, seed=0x5200 + len(kind)), kind)


@pytest.mark.parametrize("kind", ["f32", "f64"], ids=["radix_sort_f32", "radix_sort_f64"])
def test_ref_floats(kind, cuda):  # tests.rs:133-173: Uniform(-1,1) + {0.0,-0.0,NAN,INF,-INF}
    _check_host(_ref.float_test_keys(kind, N_REF, seed=77), kind)


def test_ref_tuple(cuda):  # tests.rs:175-187: random (u32,u32), stable sort_by_key(|e| e.0)
    rng = np.random.default_rng(1234)
    recs = rng.integers(0, 2**32, size=(N_REF, 2), dtype=np.uint64).astype(np.uint32)
    want = _ref.expected_sorted_records(recs, "u32")
    want_or = recs.copy()
    cpu_oracle.sort0(want_or, "pair_u32")
    assert _ref.same_bits(want, want_or)
    got = recs.copy()
    rs.radix_sort_records(got, rs.Kind.U32, key_off=0)
    assert _ref.same_bits(got, want)


# ---- every kind, device-resident path, edge sizes ----------------------------------

# tile sizes of the shipped K3 variants (csrc/sort_impl.cuh): 3072 (16-byte), 6144 (8-byte records),
# 8192 (8-byte keys), 12288 (2-byte), 13312 (4-byte), 16384 (1-byte)
EDGE_SIZES = [1, 2, 31, 32, 33, 95, 96, 97, 255, 256, 257, 3071, 3072, 3073, 6143, 6144, 6145, 8191, 8192, 8193,
              12287, 12288, 12289, 13311, 13312, 13313, 16383, 16384, 16385, 26625, 100_003]


@pytest.mark.parametrize("kind", _ref.KINDS)
def test_device_path_edge_sizes(kind, cuda):
    for n in EDGE_SIZES + [4095, 4096, 4097]:  # 4096: the largest slice the one-CTA small sort takes
        _check_device(_ref.random_keys(kind, n, seed=n), kind, cuda)


@pytest.mark.parametrize("kind", _ref.KINDS)
def test_small_sizes_through_the_big_path(kind, cuda):
    # slices of <= 4096 elements normally take the one-CTA small sort: switch it off so that
    # the K1-K2-K3 sequence keeps its coverage of ragged single tiles
    rs.debug_small_sort(False)
    try:
        for n in [s for s in EDGE_SIZES if s <= 4096] + [4095, 4096]:
            _check_device(_ref.random_keys(kind, n, seed=n), kind, cuda)
    finally:
        rs.debug_small_sort(True)


# ---- 1- and 2-byte keys: value histogram + fill (csrc/narrow_kernels.cuh; counting_sort, mod.rs:40-59) ----

NARROW = ["u8", "i8", "u16", "i16"]


@pytest.mark.parametrize("kind", NARROW)
def test_narrow_keys_counting_path(kind, cuda):
    """The default route of key-only u8/i8/u16/i16 above the small-sort size: K1n + K2n + K3n.
    Sizes around the fill kernel's chunking (4096-element granules, 65536 per CTA, 296 CTAs),
    unaligned slices, and value distributions from one value to all of them."""
    import torch
    dt = {"u8": np.uint8, "i8": np.int8, "u16": np.uint16, "i16": np.int16}[kind]
    launches0 = rs.launch_count()
    for n in (4097, 65535, 65536, 65537, 1_000_003, (1 << 22) + 5, 296 * 65536 + 4097):
        _check_device(_ref.random_keys(kind, n, seed=n), kind, cuda)
    rng = np.random.default_rng(5)
    n = 3_000_001
    cases = {
        "all_equal": np.full(n, -7 if kind[0] == "i" else 200, dtype=dt),
        "two_values": rng.choice(np.array([1, np.iinfo(dt).max], dtype=dt), size=n),
        "extremes": rng.choice(np.array([np.iinfo(dt).min, -1 if kind[0] == "i" else 1, 0, np.iinfo(dt).max], dtype=dt), size=n),
        "ascending": np.arange(n).astype(dt),
        "low_values": (rng.integers(0, 37, size=n)).astype(dt),
    }
    for name, a in cases.items():
        _check_device(a, kind, cuda)
        _check_host(a, kind)
    # slices that start 1..3 elements into an allocation (head / tail handling of the 16-byte vectors)
    base = _ref.random_keys(kind, 500_011, seed=9)
    t_all = torch.from_numpy(base).to(cuda)
    for off in (1, 2, 3, 7):
        t = t_all[off:off + 400_000]
        a = t.cpu().numpy().copy()
        want = a.copy()
        cpu_oracle.sort0(want, kind)
        rs.radix_sort(t, _kind(kind))
        torch.cuda.synchronize()
        assert _ref.same_bits(t.cpu().numpy(), want), f"offset {off}"
        # the elements around the slice are untouched
        assert _ref.same_bits(t_all[:off].cpu().numpy(), base[:off])
        base = t_all.cpu().numpy().copy()
    assert rs.launch_count() > launches0


@pytest.mark.parametrize("kind", NARROW)
def test_narrow_keys_through_the_onesweep_passes(kind, cuda):
    """The same widths with the counting path switched off: K1-K2-K3 as for every other width."""
    rs.debug_narrow_fill(False)
    try:
        for n in (4097, 100_003, 1_000_003):
            _check_device(_ref.random_keys(kind, n, seed=n + 1), kind, cuda)
        _check_device(np.full(70_000, 3, dtype={"u8": np.uint8, "i8": np.int8, "u16": np.uint16, "i16": np.int16}[kind]), kind, cuda)
    finally:
        rs.debug_narrow_fill(True)


@pytest.mark.parametrize("small", [True, False])
@pytest.mark.parametrize("key_kind,dt", [("u32", np.uint32), ("i64", np.int64), ("f32", np.float32)])
def test_small_record_slices_stay_stable(key_kind, dt, small, cuda):
    import torch
    kb = np.dtype(dt).itemsize
    udt = np.uint32 if kb == 4 else np.uint64
    rs.debug_small_sort(small)
    try:
        for n in (1, 2, 33, 257, 1000, 4095, 4096, 4097):
            rng = np.random.default_rng(n)
            recs = np.empty((n, 2), dtype=udt)
            recs[:, 0] = rng.integers(0, 2**(8 * kb), size=n, dtype=np.uint64).astype(udt) & udt(0x80000003 if kb == 4 else 0x8000000000000003)
            recs[:, 1] = np.arange(n, dtype=udt)
            want = _ref.expected_sorted_records(recs.view(dt), key_kind, 0)
            t = torch.from_numpy(recs.view(np.int32 if kb == 4 else np.int64)).to(cuda)
            rs.radix_sort_records(t, _kind(key_kind), key_off=0)
            torch.cuda.synchronize()
            assert _ref.same_bits(t.cpu().numpy(), want), f"n={n}"
    finally:
        rs.debug_small_sort(True)


@pytest.mark.parametrize("kind", _ref.KINDS)
def test_golden_fixtures(kind, cuda):
    """The committed fixtures of tests/golden/ through the C ABI: host slice and device buffer."""
    import torch
    inp, want = _ref.golden(kind)
    got = inp.copy()
    rs.radix_sort(got, _kind(kind))
    assert _ref.same_bits(got, want), "host path"
    t = _to_device(inp, cuda)
    rs.radix_sort(t, _kind(kind))
    torch.cuda.synchronize()
    assert _ref.same_bits(t.cpu().numpy(), want), "device path"


def test_golden_pairs(cuda):
    inp, want = _ref.golden_pairs()
    got = inp.copy()
    rs.radix_sort_records(got, "u32", 0)
    assert _ref.same_bits(got, want)


@pytest.mark.parametrize("kind", _ref.KINDS)
def test_host_path_small_and_odd(kind, cuda):
    for n in (1, 2, 97, 1_000_001):
        _check_host(_ref.random_keys(kind, n, seed=n + 7), kind)


@pytest.mark.parametrize("kind", _ref.KINDS)
def test_host_path_pipelined(kind, cuda):
    """The large-slice form of b200sort_keys_host (16 key ranges, copy-out overlapped
    with the range sorts), forced at small sizes: same bits as the plain path."""
    rs.set_host_pipeline_min_bytes(0)
    try:
        for n in (4096, 100_003, 1_000_001):
            _check_host(_ref.random_keys(kind, n, seed=n + 11), kind)
        # skewed inputs: empty ranges, one range holding everything, repeated splitters
        n = 200_000
        base = _ref.random_keys(kind, n, seed=5)
        allsame = base.copy(); allsame[:] = base[0]
        few = base.copy(); few[:] = base[np.arange(n) % 3]
        mostly = base.copy(); mostly[: n - 100] = base[0]
        presorted = _ref.expected_sorted(base, kind)
        for a in (allsame, few, mostly, presorted, np.ascontiguousarray(presorted[::-1])):
            _check_host(a, kind)
        if kind in ("f32", "f64"):
            _check_host(_ref.float_test_keys(kind, 150_000, seed=3), kind)
    finally:
        rs.set_host_pipeline_min_bytes(64 << 20)


def test_host_path_large_pageable_records(cuda):
    """A pageable (key, value) slice of 64 MiB or more goes through the library's
    pinned bounce buffers (host_staging.cuh) in both directions; a ragged tail that
    is not a multiple of the 16 MiB bounce buffer must arrive intact."""
    n = (80 << 20) // 8 + 12_345
    rng = np.random.default_rng(99)
    recs = rng.integers(0, 2**32, size=(n, 2), dtype=np.uint64).astype(np.uint32)
    recs[:, 0] &= 0xFFFF  # many equal keys: stability is visible in the values
    want = _ref.expected_sorted_records(recs, "u32")
    got = recs.copy()
    rs.radix_sort_records(got, rs.Kind.U32, key_off=0)
    assert _ref.same_bits(got, want)


@pytest.mark.parametrize("kind", ["u8", "u16", "u32", "u64", "i32", "i64", "f32", "f64", "u128"])
def test_degenerate_inputs_skip_and_no_skip(kind, cuda):
    import torch
    n = 300_000
    kb = _ref.KEY_BYTES[kind]
    base = _ref.random_keys(kind, n, seed=1)
    cases = {
        "all_equal": np.repeat(base[:1], n, axis=0),
        "two_values": base[np.arange(n) % 2],
        "low_byte_only": None,
    }
    flat = base.view(np.uint8).reshape(n, kb).copy()
    flat[:, 1:] = 0x5A if kb > 1 else flat[:, 1:]
    cases["low_byte_only"] = flat.view(base.dtype).reshape(base.shape)
    srt = base.copy(); cpu_oracle.sort0(srt, kind)
    cases["ascending"] = srt
    cases["descending"] = srt[::-1].copy()
    for name, a in cases.items():
        a = np.ascontiguousarray(a)
        _check_device(a, kind, cuda)
        _check_device(a, kind, cuda, flags=rs.FLAG_NO_SKIP)


@pytest.mark.parametrize("kind", ["u8", "u32", "i64", "f32", "u128"])
def test_degenerate_small_slices(kind, cuda):
    # the one-CTA small sort skips a digit on which all keys agree: with and without the skip
    for n in (1, 37, 3000, 4096):
        base = _ref.random_keys(kind, n, seed=n + 3)
        for a in (np.repeat(base[:1], n, axis=0), base[np.arange(n) % 2]):
            a = np.ascontiguousarray(a)
            _check_device(a, kind, cuda)
            _check_device(a, kind, cuda, flags=rs.FLAG_NO_SKIP)


def test_known_answer_float_order(cuda):
    # -NaN < -inf < ... < -0 < +0 < ... < +inf < +NaN, NaNs by payload
    bits = np.array([0xFFC00000, 0xFF800001, 0xFF800000, 0xBF800000, 0x80000001, 0x80000000, 0x00000000,
                     0x00000001, 0x3F800000, 0x7F800000, 0x7F800001, 0x7FC00000, 0x7FFFFFFF], dtype=np.uint32)
    a = bits[::-1].copy().view(np.float32)
    rs.radix_sort(a)
    assert a.view(np.uint32).tolist() == bits.tolist()
    b = np.array([0x7FFFFFFF, 0, 0x80000000, 0xFFFFFFFF, 1], dtype=np.uint32).view(np.int32)
    rs.radix_sort(b)
    assert b.tolist() == [-(2**31), -1, 0, 1, 2**31 - 1]


def test_f64_specials_and_subnormals(cuda):
    vals = np.array([np.inf, -0.0, 1.0, np.nan, 0.0, -np.inf, -1.0, -np.nan, 5e-324, -5e-324,
                     1.7976931348623157e308, -1.7976931348623157e308] * 1000, dtype=np.float64)
    _check_host(vals, "f64")


# ---- distribution families of distr.rs, generated on the device (K5) ---------------

@pytest.mark.parametrize("kind,dist,param", [
    ("u32", distr.Distr.UNIFORM, 0), ("u64", distr.Distr.UNIFORM, 0), ("u128", distr.Distr.UNIFORM, 0),
    ("u32", distr.Distr.GEOMETRIC, 2.0 ** -10), ("u64", distr.Distr.GEOMETRIC, 2.0 ** -10),
    ("u32", distr.Distr.ZIPF, 0.75), ("u32", distr.Distr.ZIPF, 1.0), ("u32", distr.Distr.ZIPF, 1.5),
    ("u64", distr.Distr.ZIPF, 1.0),
    ("u32", distr.Distr.STEP, 1), ("u32", distr.Distr.STEP, 2), ("u32", distr.Distr.STEP, 16),
    ("u32", distr.Distr.STEP, 256), ("u32", distr.Distr.STEP, 65536), ("u64", distr.Distr.STEP, 256),
    ("u32", distr.Distr.ASCENDING, 2_000_003), ("u32", distr.Distr.DESCENDING, 2_000_003),
    ("u64", distr.Distr.DESCENDING, 2_000_003),
    ("f32", distr.Distr.FLOAT_TEST, 1000), ("f64", distr.Distr.FLOAT_TEST, 1000),
    ("f32", distr.Distr.UNIFORM, 0), ("f64", distr.Distr.UNIFORM, 0),
    ("i32", distr.Distr.UNIFORM, 0), ("i64", distr.Distr.UNIFORM, 0),
])
def test_distribution_families(kind, dist, param, cuda):
    import torch
    n = 2_000_003
    t = distr.empty_keys(n, _kind(kind), cuda)
    distr.generate(t, _kind(kind), dist, param, seed=0xB2000004)
    torch.cuda.synchronize()
    a = t.cpu().numpy()
    if dist == distr.Distr.UNIFORM:  # the numpy mirror of the stream is bit-identical
        assert _ref.same_bits(a, distr.uniform_host(n, _kind(kind), seed=0xB2000004))
    if dist == distr.Distr.STEP:
        assert len(np.unique(a)) == min(int(param), n)
    want = a.copy()
    cpu_oracle.sort0(want, kind)
    fp_before = verify.fingerprint(t, _kind(kind))
    rs.radix_sort(t, _kind(kind))
    torch.cuda.synchronize()
    assert _ref.same_bits(t.cpu().numpy(), want)
    assert verify.count_unsorted(t, _kind(kind)) == 0
    assert verify.fingerprint(t, _kind(kind)) == fp_before


# ---- records: (K, V) tuples, stable (radix_digits.rs:126-136) ----------------------

@pytest.mark.parametrize("key_kind,dt", [("u32", np.uint32), ("i32", np.int32), ("f32", np.float32),
                                         ("u64", np.uint64), ("i64", np.int64), ("f64", np.float64)])
@pytest.mark.parametrize("key_col", [0, 1])
def test_records_stable_host_and_device(key_kind, dt, key_col, cuda):
    import torch
    rng = np.random.default_rng(99)
    n = 500_003
    kb = np.dtype(dt).itemsize
    udt = np.uint32 if kb == 4 else np.uint64
    recs = np.empty((n, 2), dtype=udt)
    # few distinct keys -> long runs of equal keys -> stability is observable
    keys = rng.integers(0, 2**(8 * kb), size=n, dtype=np.uint64).astype(udt)
    keys[rng.random(n) < 0.9] &= udt(0xFF0000FF if kb == 4 else 0xFF000000000000FF)
    recs[:, key_col] = keys
    recs[:, 1 - key_col] = np.arange(n, dtype=udt)  # payload = original position
    typed = recs.view(dt)
    want = _ref.expected_sorted_records(typed, key_kind, key_col)
    got = typed.copy()
    rs.radix_sort_records(got, _kind(key_kind), key_off=key_col * kb)
    assert _ref.same_bits(got, want)
    t = torch.from_numpy(recs.view(np.int32 if kb == 4 else np.int64)).to(cuda)
    rs.radix_sort_records(t, _kind(key_kind), key_off=key_col * kb)
    torch.cuda.synchronize()
    assert _ref.same_bits(t.cpu().numpy(), want)


# ---- the K3 slow path (ballot re-rank of a tile whose order check failed) ---------------
# It has never fired on B200 (the fast ranker's lane-order assumption holds), so tests force it.

@pytest.mark.parametrize("kind", ["u8", "u16", "u32", "i64", "f64", "u128"])
def test_forced_rerank_keys(kind, cuda):
    import torch
    k = _kind(kind)
    n = 300_007
    a = _ref.random_keys(kind, n, seed=5)
    want = _ref.expected_sorted(a, kind)
    rs.debug_force_rerank(True)
    rs.debug_narrow_fill(False)  # 1- and 2-byte keys: through K3, which is what this test is about
    try:
        for variant in range(rs.num_variants(rs.key_bytes(k))):
            rs.set_variant(rs.key_bytes(k), variant)
            t = _to_device(a, cuda)
            ws = torch.empty(rs.temp_bytes(n, k), dtype=torch.uint8, device=cuda)
            rs.radix_sort(t, k, workspace=ws)
            torch.cuda.synchronize()
            assert _ref.same_bits(t.cpu().numpy(), want), f"variant {variant}"
            assert rs.debug_reranked_tiles(ws, n, k) > 0, "the slow path did not run"
    finally:
        rs.set_variant(rs.key_bytes(k), -1)
        rs.debug_force_rerank(False)
    try:
        t = _to_device(a, cuda)
        ws = torch.empty(rs.temp_bytes(n, k), dtype=torch.uint8, device=cuda)
        rs.radix_sort(t, k, workspace=ws)
        torch.cuda.synchronize()
        assert rs.debug_reranked_tiles(ws, n, k) == 0, "a tile failed its order check on the fast path"
    finally:
        rs.debug_narrow_fill(True)


@pytest.mark.parametrize("key_kind,dt", [("u32", np.uint32), ("f64", np.float64)])
def test_forced_rerank_records_stay_stable(key_kind, dt, cuda):
    import torch
    rng = np.random.default_rng(123)
    n = 200_003
    kb = np.dtype(dt).itemsize
    udt = np.uint32 if kb == 4 else np.uint64
    recs = np.empty((n, 2), dtype=udt)
    recs[:, 0] = rng.integers(0, 2**(8 * kb), size=n, dtype=np.uint64).astype(udt) & udt(0xFF0000FF if kb == 4 else 0xFF000000000000FF)
    recs[:, 1] = np.arange(n, dtype=udt)
    want = _ref.expected_sorted_records(recs.view(dt), key_kind, 0)
    k = _kind(key_kind)
    L = _ffi.lib()
    for force in (True, False):
        rs.debug_force_rerank(force)
        try:
            t = torch.from_numpy(recs.view(np.int32 if kb == 4 else np.int64)).to(cuda)
            ws = torch.empty(int(L.b200sort_records_temp_bytes(n, 2 * kb, int(k))), dtype=torch.uint8, device=cuda)
            rs.radix_sort_records(t, k, key_off=0, workspace=ws)
            torch.cuda.synchronize()
            assert _ref.same_bits(t.cpu().numpy(), want), f"forced={force}"
            took = rs.debug_reranked_tiles(ws, n, k, rec_bytes=2 * kb)
            assert (took > 0) == force, f"forced={force}: {took} tiles took the slow path"
        finally:
            rs.debug_force_rerank(False)


def test_demo_shapes_pairs_payload_zero(cuda):
    # main.rs:101-126 + distr.rs:42-52: KeyUniform pairs (random key, payload 0)
    for udt, kind, oracle_kind in ((np.uint32, rs.Kind.U32, "pair_u32"), (np.uint64, rs.Kind.U64, "pair_u64")):
        n = 1_000_000
        recs = np.zeros((n, 2), dtype=udt)
        recs[:, 0] = distr.uniform_host(n, kind, seed=5)
        want = recs.copy()
        cpu_oracle.sort0(want, oracle_kind)
        rs.radix_sort_records(recs, kind)
        assert _ref.same_bits(recs, want)


# ---- kernel variants, alignment, re-entrancy, graphs --------------------------------

@pytest.mark.parametrize("kind,kb", [("u8", 1), ("u16", 2), ("u32", 4), ("u64", 8), ("u128", 16)])
def test_every_kernel_variant(kind, kb, cuda):
    a = _ref.random_keys(kind, 777_777, seed=3)
    try:
        for v in range(rs.num_variants(kb)):
            rs.set_variant(kb, v)
            _check_device(a, kind, cuda)
    finally:
        rs.set_variant(kb, -1)


@pytest.mark.parametrize("kind", ["u8", "u16", "u32", "u64"])
def test_unaligned_device_slice(kind, cuda):
    import torch
    a = _ref.random_keys(kind, 200_003, seed=8)
    t = _to_device(a, cuda)
    for off in (1, 3):
        sub = t[off:]
        want = a[off:].copy()
        cpu_oracle.sort0(want, kind)
        rs.radix_sort(sub, _kind(kind))
        torch.cuda.synchronize()
        assert _ref.same_bits(sub.cpu().numpy(), want)
        assert _ref.same_bits(t[:off].cpu().numpy(), a[:off])  # prefix untouched
        t = _to_device(a, cuda)


def test_concurrent_host_calls(cuda):
    # libtest runs the reference's tests on parallel threads
    kinds = ["u32", "u64", "i32", "f64", "u16", "u128", "i8", "f32"]
    inputs = [_ref.random_keys(k, 400_000 + 1000 * i, seed=i) for i, k in enumerate(kinds)]
    outs = [a.copy() for a in inputs]
    errs = []

    def work(i):
        try:
            for _ in range(3):
                outs[i][...] = inputs[i]
                rs.radix_sort(outs[i], _kind(kinds[i]))
        except Exception as exc:  # noqa: BLE001
            errs.append(exc)

    th = [threading.Thread(target=work, args=(i,)) for i in range(len(kinds))]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errs, errs
    for k, a, got in zip(kinds, inputs, outs):
        want = a.copy()
        cpu_oracle.sort0(want, k)
        assert _ref.same_bits(got, want)


def test_concurrent_host_calls_pipelined(cuda):
    """The same, every call forced through the large-slice form (chunk splits, range
    sorts, shared pinned bounce buffers and copy threads): pageable and pinned slices
    sorted from four threads at once."""
    import torch
    kinds = ["u64", "i32", "f64", "u128"]
    inputs = [_ref.random_keys(k, 700_000 + 4097 * i, seed=40 + i) for i, k in enumerate(kinds)]
    outs = [a.copy() for a in inputs]
    pinned = torch.empty(inputs[0].nbytes, dtype=torch.uint8, pin_memory=True)  # thread 0 sorts a pinned slice
    outs[0] = pinned.numpy().view(inputs[0].dtype)
    errs = []

    def work(i):
        try:
            for _ in range(3):
                outs[i][...] = inputs[i]
                rs.radix_sort(outs[i], _kind(kinds[i]))
        except Exception as exc:  # noqa: BLE001
            errs.append(exc)

    rs.set_host_pipeline_min_bytes(0)
    try:
        th = [threading.Thread(target=work, args=(i,)) for i in range(len(kinds))]
        [t.start() for t in th]
        [t.join() for t in th]
    finally:
        rs.set_host_pipeline_min_bytes(64 << 20)
    assert not errs, errs
    for k, a, got in zip(kinds, inputs, outs):
        assert _ref.same_bits(got, _ref.expected_sorted(a, k))


def test_cuda_graph_capture_of_device_sort(cuda):
    import torch
    n = 1_000_000
    a = _ref.random_keys("u32", n, seed=21)
    want = a.copy()
    cpu_oracle.sort0(want, "u32")
    src = _to_device(a, cuda)
    keys = torch.empty_like(src)
    ws = torch.empty(rs.temp_bytes(n, rs.Kind.U32), dtype=torch.uint8, device=cuda)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        keys.copy_(src)
        rs.radix_sort(keys, workspace=ws)  # warm: sets function attributes outside capture
    s.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        keys.copy_(src)
        rs.radix_sort(keys, workspace=ws)
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    assert _ref.same_bits(keys.cpu().numpy(), want)


def test_workspace_too_small_is_rejected(cuda):
    import torch
    t = torch.zeros(1000, dtype=torch.int32, device=cuda)
    with pytest.raises(rs.B200SortError) as e:
        rs.radix_sort(t, workspace=torch.empty(256, dtype=torch.uint8, device=cuda))
    assert e.value.code == 1


def test_copyonly_aliases_and_counting_sort(cuda):
    a = _ref.random_keys("u8", 100_000, seed=4)
    want = np.sort(a)
    b = a.copy(); rs.counting_sort(b)
    assert _ref.same_bits(b, want)
    for fn in (rs.radix_sort0, rs.radix_sort1, rs.radix_sort2, rs.radix_sort3, rs.radix_sort4, rs.radix_sort5):
        c = a.copy(); fn(c)
        assert _ref.same_bits(c, want)


@pytest.mark.parametrize("kind,n", [("u64", 3_000_001), ("f32", 1_000_000), ("i16", 700_000), ("u32", 5000), ("u64", 0)])
def test_pooled_workspace_entry_point(kind, n, cuda):
    """b200sort_keys_device_pooled: the temp buffer is taken and returned inside the call (the reference's own
    convention, mod.rs:71-83,176) from the library's stream-ordered pool.  Same bits as the oracle, call after call;
    b200sort_release_cached_memory hands the pool's idle memory back and the next call simply allocates again."""
    import torch
    L = _ffi.lib()
    a = _ref.random_keys(kind, n, seed=73)
    want = a.copy()
    cpu_oracle.sort0(want, kind)
    with torch.cuda.device(cuda):
        st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        for round_ in range(3):
            t = _to_device(a, cuda) if n else torch.empty(0, dtype=torch.uint64, device=cuda)
            _ffi.check(L.b200sort_keys_device_pooled(t.data_ptr() if n else None, n, int(_kind(kind)), st))
            torch.cuda.synchronize()
            assert _ref.same_bits(t.cpu().numpy(), want), f"{kind}: pooled call {round_} differs from the oracle"
            if round_ == 1:
                free_before = torch.cuda.mem_get_info()[0]
                assert L.b200sort_release_cached_memory() == 0
                if n >= 700_000:  # the pool held at least the ping-pong buffer of this sort
                    assert torch.cuda.mem_get_info()[0] > free_before


# ---- multi-GPU building block: K7 partition ------------------------------------------

@pytest.mark.parametrize("kind", ["u32", "u64", "i64", "f32"])
def test_partition_by_splitters(kind, cuda):
    import torch
    n = 1_500_001
    a = _ref.random_keys(kind, n, seed=31)
    srt = _ref.expected_sorted(a, kind)
    nsplit = 7
    spl = np.ascontiguousarray(srt[[n * (j + 1) // (nsplit + 1) for j in range(nsplit)]])
    # destination of every key in get_digit's order: its rank among the splitters
    if kind in ("f32", "f64"):
        keyv = _ref.total_cmp_key(a.view(np.int32 if kind == "f32" else np.int64))
        splv = _ref.total_cmp_key(spl.view(np.int32 if kind == "f32" else np.int64))
    else:
        keyv, splv = a, spl
    dest = np.searchsorted(splv, keyv, side="right")  # #{j : spl[j] <= key}
    want = np.concatenate([a[dest == r] for r in range(nsplit + 1)])  # stable by destination
    want_counts = np.bincount(dest, minlength=nsplit + 1)

    L = _ffi.lib()
    d_in = _to_device(a, cuda)
    d_out = torch.empty_like(d_in)
    ws = torch.empty(rs.temp_bytes(n, _kind(kind)), dtype=torch.uint8, device=cuda)
    counts = torch.zeros(nsplit + 1, dtype=torch.int64, device=cuda)
    st = torch.cuda.current_stream().cuda_stream
    rc = L.b200sort_partition_device(d_in.data_ptr(), d_out.data_ptr(), ws.data_ptr(), ws.numel(), n,
                                     int(_kind(kind)), spl.ctypes.data, nsplit, counts.data_ptr(),
                                     ctypes.c_void_p(st))
    assert rc == 0, L.b200sort_last_error()
    torch.cuda.synchronize()
    assert counts.cpu().numpy().tolist() == want_counts.tolist()
    assert _ref.same_bits(d_out.cpu().numpy(), want)
    assert _ref.same_bits(d_in.cpu().numpy(), a)  # input untouched


@pytest.mark.parametrize("kind,pieces", [("u32", 1), ("u64", 1), ("f64", 1), ("u64", 2), ("u32", 3)])
def test_partition_count_then_exchange_into_separate_buffers(kind, pieces, cuda):
    """The fused multi-GPU form of K7 on one device: every destination has its own
    receive buffer ("peer" pointer) and this rank's keys start at a given offset in
    each, as if lower ranks had already claimed the space before it."""
    import torch
    n = 1_200_003
    a = _ref.random_keys(kind, n, seed=57)
    srt = _ref.expected_sorted(a, kind)
    nsplit = 3
    spl = np.ascontiguousarray(srt[[n * (j + 1) // (nsplit + 1) for j in range(nsplit)]])
    if kind in ("f32", "f64"):
        keyv = _ref.total_cmp_key(a.view(np.int32 if kind == "f32" else np.int64))
        splv = _ref.total_cmp_key(spl.view(np.int32 if kind == "f32" else np.int64))
    else:
        keyv, splv = a, spl
    dest = np.searchsorted(splv, keyv, side="right")
    want_counts = np.bincount(dest, minlength=nsplit + 1)

    L = _ffi.lib()
    k = _kind(kind)
    d_in = _to_device(a, cuda)
    ws = torch.empty(rs.temp_bytes(n, k), dtype=torch.uint8, device=cuda)
    counts = torch.zeros(nsplit + 1, dtype=torch.int64, device=cuda)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = L.b200sort_partition_count_device(d_in.data_ptr(), ws.data_ptr(), ws.numel(), n, int(k), spl.ctypes.data, nsplit,
                                           counts.data_ptr(), st)
    assert rc == 0, L.b200sort_last_error()
    assert counts.cpu().numpy().tolist() == want_counts.tolist()

    offs = [5, 0, 1234, 77]  # where "my" keys start in each destination's buffer
    fill = 0xEE
    bufs = [torch.full(((offs[r] + int(want_counts[r]) + 9) * a.itemsize,), fill, dtype=torch.uint8, device=cuda)
            .view(d_in.dtype) for r in range(nsplit + 1)]
    bases = (ctypes.c_void_p * (nsplit + 1))(*[b.data_ptr() for b in bufs])
    d_offs = torch.tensor(offs, dtype=torch.int64, device=cuda)
    if pieces == 1:
        rc = L.b200sort_partition_exchange_device(d_in.data_ptr(), ws.data_ptr(), ws.numel(), n, int(k), spl.ctypes.data,
                                                  nsplit, bases, d_offs.data_ptr(), st)
        assert rc == 0, L.b200sort_last_error()
    else:
        # the overlapped exchange: the buckets go out in several launches over the same counts,
        # each storing only the buckets of its mask (pieces == 3: one launch stores nothing new)
        masks = {2: [0b0101, 0b1010], 3: [0b0001, 0b0110, 0b1000]}[pieces]
        for i, m in enumerate(masks):
            rc = L.b200sort_partition_exchange_part_device(d_in.data_ptr(), ws.data_ptr(), ws.numel(), n, int(k),
                                                           spl.ctypes.data, nsplit, bases, d_offs.data_ptr(), m, int(i > 0), st)
            assert rc == 0, L.b200sort_last_error()
            if i == 0:  # the buckets of the later pieces are still untouched
                torch.cuda.synchronize()
                later = [r for r in range(nsplit + 1) if not (m >> r) & 1]
                for r in later:
                    assert bool((bufs[r].view(torch.uint8) == fill).all())
    torch.cuda.synchronize()
    pad = np.frombuffer(bytes([fill]) * a.itemsize, dtype=a.dtype)[0]
    for r in range(nsplit + 1):
        got = bufs[r].cpu().numpy()
        c = int(want_counts[r])
        assert _ref.same_bits(got[offs[r]: offs[r] + c], a[dest == r]), f"destination {r}"
        # nothing outside [offset, offset + count) was touched
        assert _ref.same_bits(got[: offs[r]], np.full(offs[r], pad, dtype=a.dtype))
        assert _ref.same_bits(got[offs[r] + c:], np.full(9, pad, dtype=a.dtype))
    assert _ref.same_bits(d_in.cpu().numpy(), a)


@pytest.mark.parametrize("kind", ["u32", "u64", "f64"])
def test_partition_claim_two_writers_into_shared_regions(kind, cuda):
    """The count-free form of K7 (b200sort_partition_claim_device) on one device: two "ranks" (two shards,
    one launch each) claim room in the SAME four regions through the same four cursors, as two GPUs would in
    each other's receive buffers.  The order inside a region is free, so regions are compared as multisets
    (sorted) with what b200sort_partition_device gives for the concatenated input; the cursors must end at
    the bucket sizes and no overflow bit may be set."""
    import torch
    n = 1_200_003
    a = _ref.random_keys(kind, n, seed=61)
    srt = _ref.expected_sorted(a, kind)
    nsplit = 3
    spl = np.ascontiguousarray(srt[[n * (j + 1) // (nsplit + 1) for j in range(nsplit)]])
    L = _ffi.lib()
    k = _kind(kind)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    d_all = _to_device(a, cuda)
    # the reference split: the counting partition of the whole input
    d_ref = torch.empty_like(d_all)
    ws = torch.empty(rs.temp_bytes(n, k), dtype=torch.uint8, device=cuda)
    counts = torch.zeros(nsplit + 1, dtype=torch.int64, device=cuda)
    rc = L.b200sort_partition_device(d_all.data_ptr(), d_ref.data_ptr(), ws.data_ptr(), ws.numel(), n, int(k),
                                     spl.ctypes.data, nsplit, counts.data_ptr(), st)
    assert rc == 0, L.b200sort_last_error()
    cnt = counts.cpu().numpy().tolist()
    ref = d_ref.cpu().numpy()
    # two writers: shards of unequal size
    cut = 700_001
    shards = [d_all[:cut], d_all[cut:]]
    cap = [c + 17 for c in cnt]
    fill = 0xEE
    regions = [torch.full((cap[r] * a.itemsize,), fill, dtype=torch.uint8, device=cuda).view(d_all.dtype) for r in range(nsplit + 1)]
    ctrl = torch.zeros(nsplit + 2, dtype=torch.int64, device=cuda)  # cursors 0..nsplit, then the flag word
    bases = (ctypes.c_void_p * (nsplit + 1))(*[r_.data_ptr() for r_ in regions])
    cur = (ctypes.c_uint64 * (nsplit + 1))(*[ctrl.data_ptr() + 8 * r for r in range(nsplit + 1)])
    caps = (ctypes.c_uint64 * (nsplit + 1))(*cap)
    flags = ctypes.c_void_p(ctrl.data_ptr() + 8 * (nsplit + 1))
    for sh in shards:
        rc = L.b200sort_partition_claim_device(sh.data_ptr(), ws.data_ptr(), ws.numel(), int(sh.shape[0]), int(k),
                                               spl.ctypes.data, nsplit, bases, cur, caps, flags, st)
        assert rc == 0, L.b200sort_last_error()
    torch.cuda.synchronize()
    c = ctrl.cpu().numpy()
    assert c[: nsplit + 1].tolist() == cnt, "cursors must end at the bucket sizes"
    assert int(c[nsplit + 1]) == 0, "no region overflowed"
    pad = np.frombuffer(bytes([fill]) * a.itemsize, dtype=a.dtype)[0]
    off = 0
    for r in range(nsplit + 1):
        got = regions[r].cpu().numpy()
        want = ref[off: off + cnt[r]]
        off += cnt[r]
        assert _ref.same_bits(_ref.expected_sorted(got[: cnt[r]].copy(), kind), _ref.expected_sorted(want.copy(), kind)), f"region {r}"
        assert _ref.same_bits(got[cnt[r]:], np.full(cap[r] - cnt[r], pad, dtype=a.dtype)), f"region {r}: stores past the claim"
    assert _ref.same_bits(d_all.cpu().numpy(), a)  # input untouched


def test_partition_claim_overflow_sets_the_flag_and_keeps_counting(cuda):
    """A region that is too small: the tiles whose claim passes the capacity store nothing and set the bucket's
    bit in the flag word; the cursor still counts every key, so the caller learns the true size and can take
    the exact (counting) path instead."""
    import torch
    kind, n, nsplit = "u64", 900_000, 3
    a = _ref.random_keys(kind, n, seed=62)
    srt = _ref.expected_sorted(a, kind)
    spl = np.ascontiguousarray(srt[[n * (j + 1) // (nsplit + 1) for j in range(nsplit)]])
    dest = np.searchsorted(spl, a, side="right")
    cnt = np.bincount(dest, minlength=nsplit + 1).tolist()
    L = _ffi.lib()
    k = _kind(kind)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    d_in = _to_device(a, cuda)
    ws = torch.empty(rs.temp_bytes(n, k), dtype=torch.uint8, device=cuda)
    cap = list(cnt)
    cap[2] = cnt[2] // 2  # bucket 2 does not fit
    fill = 0xEE
    regions = [torch.full(((cap[r] + 8) * 8,), fill, dtype=torch.uint8, device=cuda).view(d_in.dtype) for r in range(nsplit + 1)]
    ctrl = torch.zeros(nsplit + 2, dtype=torch.int64, device=cuda)
    bases = (ctypes.c_void_p * (nsplit + 1))(*[r_.data_ptr() for r_ in regions])
    cur = (ctypes.c_uint64 * (nsplit + 1))(*[ctrl.data_ptr() + 8 * r for r in range(nsplit + 1)])
    caps = (ctypes.c_uint64 * (nsplit + 1))(*cap)
    flags = ctypes.c_void_p(ctrl.data_ptr() + 8 * (nsplit + 1))
    rc = L.b200sort_partition_claim_device(d_in.data_ptr(), ws.data_ptr(), ws.numel(), n, int(k), spl.ctypes.data, nsplit,
                                           bases, cur, caps, flags, st)
    assert rc == 0, L.b200sort_last_error()
    torch.cuda.synchronize()
    c = ctrl.cpu().numpy()
    assert c[: nsplit + 1].tolist() == cnt, "the cursors hold the true bucket sizes even for the bucket that overflowed"
    assert int(c[nsplit + 1]) == 0b0100, "exactly bucket 2 is flagged"
    pad = np.frombuffer(bytes([fill]) * 8, dtype=a.dtype)[0]
    for r in range(nsplit + 1):
        got = regions[r].cpu().numpy()
        assert _ref.same_bits(got[cap[r]:], np.full(8, pad, dtype=a.dtype)), f"region {r}: a store went past its capacity"
        if r != 2:
            assert _ref.same_bits(np.sort(got[: cnt[r]]), np.sort(a[dest == r])), f"region {r}"


# ---- full-size runs: size-independent properties (sorted + same multiset) ------------

def _property_run(kind, n, dist, param, cuda, seed=0xB2000002, flags=0):
    import torch
    k = _kind(kind)
    t = distr.empty_keys(n, k, cuda)
    distr.generate(t, k, dist, param, seed=seed)
    fp = verify.fingerprint(t, k)
    assert fp[0] == n
    ws = torch.empty(rs.temp_bytes(n, k), dtype=torch.uint8, device=cuda)
    rs.radix_sort(t, k, workspace=ws, flags=flags)
    torch.cuda.synchronize()
    assert verify.count_unsorted(t, k) == 0, "not sorted"
    assert verify.fingerprint(t, k) == fp, "multiset changed"
    del ws, t
    torch.cuda.empty_cache()


def test_config2_1b_u64_properties(cuda):  # BASELINE.json configs[1]
    _property_run("u64", 1_000_000_000, distr.Distr.UNIFORM, 0, cuda)


@pytest.mark.parametrize("kind,dist,param", [("f32", distr.Distr.FLOAT_TEST, 1 << 20), ("f64", distr.Distr.UNIFORM, 0),
                                             ("i32", distr.Distr.UNIFORM, 0), ("i64", distr.Distr.UNIFORM, 0)])
def test_config3_transform_kinds_properties(kind, dist, param, cuda):  # configs[2] at 2^28 keys
    _property_run(kind, 1 << 28, dist, param, cuda)


@pytest.mark.parametrize("kind,dist,param", [("f32", distr.Distr.FLOAT_TEST, 0.0), ("f32", distr.Distr.UNIFORM, 0),
                                             ("f64", distr.Distr.UNIFORM, 0), ("i32", distr.Distr.UNIFORM, 0),
                                             ("i64", distr.Distr.UNIFORM, 0)],
                         ids=["f32_test_shaped", "f32_raw_bits", "f64_raw_bits", "i32", "i64"])
def test_config3_at_its_stated_size_1b_keys(kind, dist, param, cuda):
    """BASELINE.json configs[2] as stated: 1 000 000 000 keys of each transform kind (radix_digits.rs:55-124), the
    f32 case on both datasets bench.py times (tests.rs:135-143 shaped: Uniform[-1,1) + specials; and raw bit
    patterns: every NaN payload, both zeros, subnormals).  Sorted in get_digit's order + same multiset."""
    _property_run(kind, 1_000_000_000, dist, param, cuda, seed=0xB2000003)


@pytest.mark.parametrize("dist,param", [(distr.Distr.ZIPF, 1.0), (distr.Distr.STEP, 16), (distr.Distr.ASCENDING, 2_000_000_000)])
def test_config4_2b_u32_skew_properties(dist, param, cuda):  # configs[3]: n > 2^30 -> two look-back portions
    _property_run("u32", 2_000_000_000, dist, param, cuda)


@pytest.mark.parametrize("kind,n,m", [("u64", 1 << 29, 1), ("i64", 1 << 28, 2), ("u128", 1 << 27, 1), ("u32", 1 << 30, 3)])
def test_few_values_at_scale_no_skip(kind, n, m, cuda):
    # m distinct values put every increment of a K1 CTA on m shared-memory counters (32-bit:
    # they must not wrap below 2^32 / CTAs keys) and every K3 tile on m digits per pass;
    # NO_SKIP: the passes really run and use the bucket bases.
    _property_run(kind, n, distr.Distr.STEP, m, cuda, flags=rs.FLAG_NO_SKIP)


def test_portion_boundary_small_keys(cuda):
    # n just over one look-back portion (2^30 elements) with 1-byte keys: cheap
    # exercise of the bins_out hand-off between portions, checked exactly
    import torch
    n = (1 << 30) + 12_345
    t = torch.empty(n, dtype=torch.uint8, device=cuda)
    distr.generate(t, rs.Kind.U8, distr.Distr.UNIFORM, 0, seed=9)
    counts = torch.zeros(256, dtype=torch.int64, device=cuda)
    step = 1 << 28
    for lo in range(0, n, step):
        counts += torch.bincount(t[lo:lo + step].to(torch.int32), minlength=256)
    rs.radix_sort(t)
    torch.cuda.synchronize()
    assert verify.count_unsorted(t, rs.Kind.U8) == 0
    bounds = torch.cumsum(counts, 0).cpu().numpy()
    got = t[torch.from_numpy(np.minimum(bounds - 1, n - 1)).to(cuda)].cpu().numpy()
    present = counts.cpu().numpy() > 0
    assert (got[present] == np.arange(256)[present]).all()


# ---- all GPUs of one process behind one call (b200sort_mgpu_*) -------------------
def _check_communicator(comm, shards_np, kind):
    import torch
    from radix_sort_b200 import mgpu
    shards = [torch.from_numpy(a).to(f"cuda:{d}") for a, d in zip(shards_np, comm.devices)]
    outs = comm.sort(shards, _kind(kind))
    for t, a in zip(shards, shards_np):  # inputs are read, not written
        assert _ref.same_bits(t.cpu().numpy(), a)
    got = [o.cpu().numpy() for o in outs]
    for o, d in zip(outs, comm.devices):
        assert o.device.index == d
    whole = np.concatenate(shards_np)
    want = whole.copy()
    cpu_oracle.sort0(want, kind)
    assert _ref.same_bits(np.concatenate(got), want)  # rank order = global order
    return [len(g) for g in got]


@pytest.mark.parametrize("kind", ["u32", "u64", "i64", "f32", "f64", "u128", "i8", "u16"])
def test_communicator_shards_on_one_device(kind, cuda):
    """Three shards of one process sorted as one key set (the same device named three
    times: the count / offset / scatter / local-sort logic without needing three GPUs)."""
    from radix_sort_b200 import mgpu
    comm = mgpu.Communicator([0, 0, 0])
    try:
        sizes = _check_communicator(comm, [_ref.random_keys(kind, n, seed=70 + i) for i, n in enumerate((200_003, 150_000, 99_999))], kind)
        if kind not in ("i8", "u16"):  # few distinct values: whole values land on one shard
            assert max(sizes) < 1.2 * sum(sizes) / 3
        # ragged: an empty shard, a one-key shard; buffers of the first sort are reused
        _check_communicator(comm, [_ref.random_keys(kind, n, seed=80 + i) for i, n in enumerate((0, 1, 50_000))], kind)
        # all keys equal: everything lands on one shard
        same = _ref.random_keys(kind, 30_000, seed=3)
        same[:] = same[0]
        sizes = _check_communicator(comm, [same[:10_000].copy(), same[10_000:20_000].copy(), same[20_000:].copy()], kind)
        assert sorted(sizes) == [0, 0, 30_000]
        # nothing at all
        empty = _ref.random_keys(kind, 0, seed=1)
        assert _check_communicator(comm, [empty, empty.copy(), empty.copy()], kind) == [0, 0, 0]
    finally:
        comm.close()


def test_communicator_single_device_and_floats(cuda):
    from radix_sort_b200 import mgpu
    comm = mgpu.Communicator([0])
    try:
        _check_communicator(comm, [_ref.random_keys("u64", 300_001, seed=9)], "u64")
        _check_communicator(comm, [_ref.float_test_keys("f32", 100_000, seed=4)], "f32")
    finally:
        comm.close()
    comm = mgpu.Communicator([0, 0])
    try:
        a = _ref.float_test_keys("f64", 120_000, seed=6)  # specials: +-0, +-inf, NaNs of both signs
        _check_communicator(comm, [a[:50_000].copy(), a[50_000:].copy()], "f64")
    finally:
        comm.close()


def test_communicator_rejects_bad_arguments(cuda):
    import torch
    from radix_sort_b200 import mgpu
    with pytest.raises(rs.B200SortError):
        mgpu.Communicator([0] * 17)
    with pytest.raises(rs.B200SortError):
        mgpu.Communicator([torch.cuda.device_count() + 3])
    comm = mgpu.Communicator([0, 0])
    try:
        with pytest.raises(rs.B200SortError):
            comm.sort([torch.zeros(4, dtype=torch.int32, device=cuda)], rs.Kind.I32)  # one shard for two devices
    finally:
        comm.close()


def test_communicator_all_visible_devices(cuda):
    """One shard per GPU of the box, every key crossing NVLink to its owner (needs >= 2 GPUs)."""
    import torch
    from radix_sort_b200 import mgpu
    g = torch.cuda.device_count()
    if g < 2:
        pytest.skip("one GPU visible")
    comm = mgpu.Communicator(list(range(g)))
    try:
        for kind in ("u64", "f32", "u128"):
            _check_communicator(comm, [_ref.random_keys(kind, 400_000 + 4099 * r, seed=90 + r) for r in range(g)], kind)
    finally:
        comm.close()


@pytest.mark.parametrize("kind", ["u64", "i32", "f64", "u128", "u8"])
def test_communicator_host_slice(kind, cuda):
    """b200sort_mgpu_keys_host: one pinned host slice, every device takes a part (the same
    GPU named three times here), result in place and bit-identical to the oracle."""
    import torch
    from radix_sort_b200 import mgpu

    def pinned_copy(a):
        buf = torch.empty(max(a.nbytes, 1), dtype=torch.uint8, pin_memory=True)
        view = buf.numpy()[: a.nbytes].view(a.dtype).reshape(a.shape)
        view[...] = a
        return buf, view

    comm = mgpu.Communicator([0, 0, 0])
    L = _ffi.lib()
    L.b200sort_mgpu_set_host_min_bytes(0)
    try:
        for n in (1, 2, 5, 100_003, 600_000):
            a = _ref.random_keys(kind, n, seed=n + 17)
            keep, got = pinned_copy(a)
            comm.sort_host(got, _kind(kind))
            want = a.copy()
            cpu_oracle.sort0(want, kind)
            assert _ref.same_bits(got, want)
        comm.sort_host(_ref.random_keys(kind, 0, seed=1), _kind(kind))  # empty slice: returns
        # a pageable slice, and any slice below the threshold, is b200sort_keys_host on the first device
        a = _ref.random_keys(kind, 250_000, seed=5)
        got = a.copy()
        comm.sort_host(got, _kind(kind))
        assert _ref.same_bits(got, _ref.expected_sorted(a, kind))
        L.b200sort_mgpu_set_host_min_bytes(32 << 20)
        keep, got = pinned_copy(a)
        comm.sort_host(got, _kind(kind))
        assert _ref.same_bits(got, _ref.expected_sorted(a, kind))
    finally:
        L.b200sort_mgpu_set_host_min_bytes(32 << 20)
        comm.close()


def test_host_path_on_every_device(cuda):
    """b200sort_keys_host runs on the calling thread's current device: pooled blocks,
    cached streams and pinned bounce buffers (with their events) are per device
    (needs >= 2 GPUs)."""
    import torch
    g = torch.cuda.device_count()
    if g < 2:
        pytest.skip("one GPU visible")
    rs.set_host_pipeline_min_bytes(0)  # the pipelined form stages a pageable slice at any size
    try:
        for d in list(range(g)) + [0]:
            with torch.cuda.device(d):
                for kind in ("u64", "f32"):
                    _check_host(_ref.random_keys(kind, 300_000 + d, seed=d), kind)
        rs.set_host_pipeline_min_bytes(64 << 20)
        for d in range(g):
            with torch.cuda.device(d):
                _check_host(_ref.random_keys("i64", 200_000 + d, seed=d), "i64")
    finally:
        rs.set_host_pipeline_min_bytes(64 << 20)
