"""Generator of the committed fixtures in this directory.

The reference (Rust) cannot be built or run here and holds no golden vectors of
its own (every input of tests.rs is unseeded thread_rng), so these fixtures are
DERIVED, not taken from the reference: inputs are fixed below, expected outputs
come from the order the reference's tests assert -- slice::sort for integers
(tests.rs:12,18), sort_by(total_cmp) for floats (tests.rs:146,167), stable
sort_by_key for tuples (tests.rs:183) -- stated with Python's own big-int
sorted(), independently of oracle/ and of the CUDA code.  Run from the repo root:

    python tests/golden/make_golden.py

Both the oracle (tests/test_oracle.py) and the CUDA path (tests/test_parity_gpu.py)
are checked against the files it writes.
"""
import json
import os
import struct

HERE = os.path.dirname(os.path.abspath(__file__))
KINDS = {"u8": 1, "u16": 2, "u32": 4, "u64": 8, "u128": 16, "i8": 1, "i16": 2, "i32": 4, "i64": 8, "i128": 16, "f32": 4, "f64": 8}


def mix64(z):
    z = (z + 0x9E3779B97F4A7C15) & (2**64 - 1)
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & (2**64 - 1)
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & (2**64 - 1)
    return z ^ (z >> 31)


def order_key(bits, kind):
    """Sort key of a raw bit pattern under the reference's order (plain Python ints)."""
    w = 8 * KINDS[kind]
    if kind[0] == "u":
        return bits
    if kind[0] == "i":
        return bits - (1 << w) if bits >> (w - 1) else bits          # two's complement value: Ord
    # total_cmp (core): flip the magnitude bits of negatives, then compare as signed
    if bits >> (w - 1):
        bits ^= (1 << (w - 1)) - 1
        return bits - (1 << w)
    return bits


def specials(kind):
    w = 8 * KINDS[kind]
    top = 1 << (w - 1)
    out = [0, 1, 2, top - 1, top, top + 1, (1 << w) - 1, (1 << w) - 2, 0x55 * ((1 << w) // 255), 0xAA * ((1 << w) // 255)]
    if kind == "f32":
        out += [struct.unpack("<I", struct.pack("<f", v))[0] for v in (0.0, -0.0, 1.0, -1.0, float("inf"), float("-inf"))]
        out += [0x7FC00000, 0xFFC00000, 0x7F800001, 0xFF800001, 0x00000001, 0x80000001, 0x7FFFFFFF]   # NaNs, subnormals
    if kind == "f64":
        out += [struct.unpack("<Q", struct.pack("<d", v))[0] for v in (0.0, -0.0, 1.0, -1.0, float("inf"), float("-inf"))]
        out += [0x7FF8000000000000, 0xFFF8000000000000, 0x7FF0000000000001, 0xFFF0000000000001, 1, (1 << 63) | 1]
    return [v & ((1 << w) - 1) for v in out]


def main():
    for kind, kb in KINDS.items():
        w = 8 * kb
        n = 1031  # spans several warps, not a multiple of anything
        vals = specials(kind)
        i = 0
        while len(vals) < n:
            hi, lo = mix64(0xB200 + 7 * i), mix64(0xC0FFEE + 13 * i)
            v = ((hi << 64) | lo) & ((1 << w) - 1)
            if i % 5 == 0:
                v &= (1 << (w // 2)) - 1           # small magnitudes: constant upper digits in places
            if i % 7 == 0 and vals:
                v = vals[(i * 31) % len(vals)]     # duplicates
            vals.append(v)
            i += 1
        # deterministic shuffle so the specials are not at the front
        order = sorted(range(n), key=lambda j: mix64(j + 99))
        inp = [vals[j] for j in order]
        exp = sorted(inp, key=lambda b: order_key(b, kind))   # Python's sort is stable; equal keys are bit-identical anyway
        fmt = f"0{2 * kb}x"
        with open(os.path.join(HERE, f"{kind}.json"), "w") as fh:
            json.dump({"kind": kind, "key_bytes": kb, "n": n, "note": "derived fixture, see make_golden.py",
                       "input_hex": [format(v, fmt) for v in inp], "sorted_hex": [format(v, fmt) for v in exp]}, fh)
    # tuples (tests.rs:175-187): (u32 key with many duplicates, u32 payload = original index)
    n = 1500
    recs = [((mix64(j) % 97) * 0x01010101 & 0xFFFFFFFF, j) for j in range(n)]
    exp = sorted(recs, key=lambda r: r[0])
    with open(os.path.join(HERE, "pair_u32.json"), "w") as fh:
        json.dump({"kind": "pair_u32", "n": n, "note": "derived fixture: stable sort_by_key(|e| e.0)",
                   "input": recs, "sorted": exp}, fh)


if __name__ == "__main__":
    main()
