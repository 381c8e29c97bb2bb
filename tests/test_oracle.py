"""CPU: pin the oracle (oracle/oracle.c) to the property the reference's own
tests assert -- output == std sort of the same input -- and to the derived
known-answer vectors of SURVEY.md 8(c).  No GPU."""
import numpy as np
import pytest

from oracle import cpu_oracle
from tests import _ref

EDGE_SIZES = [0, 1, 2, 31, 32, 33, 95, 96, 97, 255, 256, 257, 1000, 4097]


@pytest.mark.parametrize("kind", _ref.KINDS)
def test_sort0_equals_std_sort_1e5(kind):
    a = _ref.random_keys(kind, 100_000, seed=11)
    want = _ref.expected_sorted(a, kind)
    cpu_oracle.sort0(a, kind)
    assert _ref.same_bits(a, want)


@pytest.mark.parametrize("kind", ["u32", "i64", "f32", "f64", "u128", "u8"])
def test_sort0_reference_size_1e6(kind):
    # the reference's tests use 1e6 elements (tests.rs:27,36,...)
    a = _ref.float_test_keys(kind, 1_000_000, 5) if kind in ("f32", "f64") else _ref.random_keys(kind, 1_000_000, 5)
    want = _ref.expected_sorted(a, kind)
    cpu_oracle.sort0(a, kind)
    assert _ref.same_bits(a, want)


@pytest.mark.parametrize("kind", _ref.KINDS)
@pytest.mark.parametrize("n", EDGE_SIZES)
def test_edge_sizes_all_variants(kind, n):
    a = _ref.random_keys(kind, n, seed=n + 1)
    want = _ref.expected_sorted(a, kind)
    for fn in (cpu_oracle.sort0, lambda b, k: cpu_oracle.sort_mt(b, k, 4), lambda b, k: cpu_oracle.sort4_mt(b, k, 4)):
        b = a.copy()
        fn(b, kind)
        assert _ref.same_bits(b, want)


@pytest.mark.parametrize("kind", ["u16", "u32", "u64", "i32", "f64", "i128"])
@pytest.mark.parametrize("threads", [1, 3, 8])
def test_threaded_variants_match_sequential(kind, threads):
    a = _ref.random_keys(kind, 300_001, seed=threads)
    want = a.copy()
    cpu_oracle.sort0(want, kind)
    b = a.copy(); cpu_oracle.sort_mt(b, kind, threads)
    c = a.copy(); cpu_oracle.sort4_mt(c, kind, threads)
    assert _ref.same_bits(b, want) and _ref.same_bits(c, want)


def test_pairs_are_stable():
    # tests.rs:175-187: random key AND payload, compared with the stable sort_by_key
    rng = np.random.default_rng(3)
    for kind, dt in (("pair_u32", np.uint32), ("pair_u64", np.uint64)):
        recs = rng.integers(0, 1 << 10, size=(200_000, 2)).astype(dt)  # many duplicate keys
        recs[:, 1] = np.arange(len(recs))
        want = _ref.expected_sorted_records(recs, "u32")
        for fn in (cpu_oracle.sort0, cpu_oracle.sort_mt, cpu_oracle.sort4_mt):
            b = recs.copy()
            fn(b, kind)
            assert _ref.same_bits(b, want)


@pytest.mark.parametrize("kind", _ref.KINDS)
def test_oracle_matches_committed_golden_fixtures(kind):
    """tests/golden/<kind>.json: fixed inputs (specials, duplicates, small magnitudes) and the
    order the reference's tests assert, stated with Python big ints (make_golden.py)."""
    inp, want = _ref.golden(kind)
    assert _ref.same_bits(_ref.expected_sorted(inp, kind), want), "numpy statement disagrees with the fixture"
    for fn in (cpu_oracle.sort0, lambda b, k: cpu_oracle.sort_mt(b, k, 3), lambda b, k: cpu_oracle.sort4_mt(b, k, 3)):
        b = inp.copy()
        fn(b, kind)
        assert _ref.same_bits(b, want)


def test_oracle_matches_golden_pairs():
    inp, want = _ref.golden_pairs()
    b = inp.copy()
    cpu_oracle.sort0(b, "pair_u32")
    assert _ref.same_bits(b, want)


F32_ORDER = [  # required ascending order, (bits, ordered key)
    (0xFFC00000, 0x003FFFFF), (0xFF800001, 0x007FFFFE), (0xFF800000, 0x007FFFFF), (0xBF800000, 0x407FFFFF),
    (0x80000001, 0x7FFFFFFE), (0x80000000, 0x7FFFFFFF), (0x00000000, 0x80000000), (0x00000001, 0x80000001),
    (0x3F800000, 0xBF800000), (0x7F800000, 0xFF800000), (0x7F800001, 0xFF800001), (0x7FC00000, 0xFFC00000),
    (0x7FFFFFFF, 0xFFFFFFFF),
]
I32_ORDER = [(0x80000000, 0x00000000), (0xFFFFFFFF, 0x7FFFFFFF), (0x00000000, 0x80000000),
             (0x00000001, 0x80000001), (0x7FFFFFFF, 0xFFFFFFFF)]


@pytest.mark.parametrize("kind,table", [("f32", F32_ORDER), ("i32", I32_ORDER)])
def test_known_answer_digits(kind, table):
    for bits, key in table:
        raw = int(bits).to_bytes(4, "little")
        got = sum(cpu_oracle.get_digit(raw, kind, i) << (8 * i) for i in range(4))
        assert got == key, hex(bits)
    # and the sort puts them in exactly this order, from a shuffled start
    want = np.array([b for b, _ in table], dtype=np.uint32)
    a = want[::-1].copy()
    cpu_oracle.sort0(a, kind)
    assert _ref.same_bits(a, want)


def test_f64_specials_order():
    vals = np.array([np.inf, -0.0, 1.0, np.nan, 0.0, -np.inf, -1.0, -np.nan, 5e-324, -5e-324], dtype=np.float64)
    want = _ref.expected_sorted(vals, "f64")
    a = vals.copy()
    cpu_oracle.sort0(a, "f64")
    assert _ref.same_bits(a, want)
    assert np.signbit(a[0]) and np.isnan(a[0]) and np.isnan(a[-1]) and not np.signbit(a[-1])


@pytest.mark.parametrize("kind", ["u32", "u64", "i16"])
def test_degenerate_inputs(kind):
    n = 10_000
    dt = _ref.NP_DTYPE[kind]
    for a in (np.full(n, 7, dtype=dt), np.arange(n).astype(dt), np.arange(n)[::-1].astype(dt),
              (np.arange(n) % 256).astype(dt)):
        want = _ref.expected_sorted(a, kind)
        b = a.copy()
        cpu_oracle.sort_mt(b, kind, 5)
        assert _ref.same_bits(b, want)
