"""Independent numpy statement of what the reference's tests compare against
(tests.rs:7-23 `slice::sort`, :146,167 `sort_by(total_cmp)`, :183 stable
`sort_by_key`).  Deliberately NOT built from get_digit: floats use the
total_cmp formulation of core (`left ^= ((left >> 63) as u64 >> 1) as i64`,
compare as signed), so it cross-checks the oracle's transform."""
import numpy as np

KINDS = ["u8", "u16", "u32", "u64", "u128", "i8", "i16", "i32", "i64", "i128", "f32", "f64"]
KIND_ID = {k: i for i, k in enumerate(KINDS)}
KEY_BYTES = {"u8": 1, "u16": 2, "u32": 4, "u64": 8, "u128": 16, "i8": 1, "i16": 2, "i32": 4, "i64": 8,
             "i128": 16, "f32": 4, "f64": 8}
NP_DTYPE = {"u8": np.uint8, "u16": np.uint16, "u32": np.uint32, "u64": np.uint64, "i8": np.int8,
            "i16": np.int16, "i32": np.int32, "i64": np.int64, "f32": np.float32, "f64": np.float64,
            "u128": np.uint64, "i128": np.uint64}


def random_keys(kind, n, seed):
    """Full-range random bit patterns, like rng.fill (tests.rs:28): (n,) array, (n, 2) uint64 for 128-bit."""
    rng = np.random.default_rng(seed)
    kb = KEY_BYTES[kind]
    raw = rng.integers(0, 256, size=n * kb, dtype=np.uint8)
    if kb == 16:
        return raw.view(np.uint64).reshape(n, 2).copy()
    return raw.view(NP_DTYPE[kind]).copy()


def float_test_keys(kind, n, seed):
    """Uniform[-1,1) with the five specials of tests.rs:139-143 dropped at random positions."""
    rng = np.random.default_rng(seed)
    dt = NP_DTYPE[kind]
    a = rng.uniform(-1.0, 1.0, size=n).astype(dt)
    pos = rng.choice(n, size=min(5, n), replace=False)
    for p, v in zip(pos, [0.0, -0.0, np.nan, np.inf, -np.inf]):
        a[p] = v
    return a


def total_cmp_key(bits_signed):
    """core::f32/f64::total_cmp key: flip the magnitude bits of negatives, compare as signed."""
    w = bits_signed.dtype.itemsize * 8
    u = bits_signed.view(np.dtype(f"uint{w}"))
    s = bits_signed >> (w - 1)  # arithmetic: 0 or -1
    mask = (s.view(u.dtype) >> np.array(1, dtype=u.dtype))
    return (u ^ mask).view(bits_signed.dtype)


def expected_sorted(a, kind):
    """What std sort gives for `a` (copy), bitwise."""
    if kind in ("u8", "u16", "u32", "u64", "i8", "i16", "i32", "i64"):
        return np.sort(a, kind="stable")
    if kind in ("f32", "f64"):
        si = a.view(np.int32 if kind == "f32" else np.int64)
        order = np.argsort(total_cmp_key(si), kind="stable")
        return a[order]
    lo, hi = a[:, 0], a[:, 1]
    if kind == "i128":
        hi = hi.view(np.int64)
    order = np.lexsort((lo, hi))
    return a[order]


def expected_sorted_records(recs, key_kind, key_col=0):
    """Stable sort_by_key(|e| e.<key_col>) of an (n, 2) array (tests.rs:183)."""
    key = recs[:, key_col]
    if key_kind in ("f32", "f64"):
        key = total_cmp_key(key.view(np.int32 if key_kind == "f32" else np.int64))
    order = np.argsort(key, kind="stable")
    return recs[order]


def same_bits(a, b):
    return a.shape == b.shape and a.dtype.itemsize == b.dtype.itemsize and a.tobytes() == b.tobytes()
