"""File-based input mode of the reference's bench driver, behind the same names.

    gen_data(size, file, distr)          main.rs:83-99   -> gen_data
    bench_sorts_data(files, sorts)       main.rs:47-81   -> bench_sorts_data

The file format is the reference's: the key array dumped as raw native-endian
bytes, no header; its length divided by the element size is the key count
(main.rs:59-60).  `size` is in GB of key bytes, `(size * 1e9 / size_of::<T>())`
elements (main.rs:88).  The reference draws from an unseeded `thread_rng`; here
the keys are the counter-based UNIFORM stream of `distr.py`, so a dump is
reproducible.  Sorting goes through the host drop-in path (`radix_sort` on a
numpy array = `b200sort_keys_host`), which is what a Rust caller of
`bench_sorts_data` would time.
"""
from __future__ import annotations

import os
import time
from typing import Callable, Iterable, Sequence

import numpy as np

from . import distr as _distr
from .api import Kind, NUMBER_OF_DIGITS, radix_sort

_CHUNK = 1 << 24  # keys generated per write: bounds host memory for multi-GB dumps


def gen_data(size: float, file, kind=Kind.U32, seed: int = 0xB2000001) -> int:
    """Write `size` GB of uniform keys of `kind` to `file` (raw, native-endian,
    main.rs:83-99).  Returns the number of keys written."""
    k = Kind(kind)
    n = int(np.float32(size) * np.float32(1e9) / np.float32(NUMBER_OF_DIGITS[k]))  # f32 arithmetic, main.rs:88
    with open(file, "wb") as fh:
        for lo in range(0, n, _CHUNK):
            _distr.uniform_host(min(_CHUNK, n - lo), k, seed=seed, index_base=lo).tofile(fh)
    return n


def load_data(file, kind) -> np.ndarray:
    """The whole file as one key array (main.rs:55-62): trailing bytes that do not
    fill an element are dropped, as `data.len() / size_of::<T>()` does."""
    k = Kind(kind)
    kb = NUMBER_OF_DIGITS[k]
    n = os.path.getsize(file) // kb
    if kb == 16:
        return np.fromfile(file, dtype=np.uint64, count=2 * n).reshape(n, 2)
    return np.fromfile(file, dtype=np.dtype(_distr._TORCH_DTYPE[k]), count=n)


def bench_sorts_data(files: Iterable, kind, sorts: Sequence[Callable] = (radix_sort,)) -> list[float]:
    """Mean seconds per file for each sort (main.rs:47-81).  With one sort the data
    is sorted in place; with several every sort gets its own copy (main.rs:63-74)."""
    files = list(files)
    if not sorts or not files:
        return []
    k = Kind(kind)
    total = [0.0] * len(sorts)
    for f in files:
        data = load_data(f, k)
        for j, sort in enumerate(sorts):
            work = data if len(sorts) == 1 else data.copy()
            tick = time.perf_counter()
            sort(work, k)
            total[j] += time.perf_counter() - tick
    return [t / len(files) for t in total]


def _main(argv=None) -> int:
    """`python -m radix_sort_b200.filebench gen <GB> <file> [kind]` /
    `... bench <kind> <file> [<file> ...]` -- the two halves of the reference's
    file-based bench (main.rs:47-99)."""
    import argparse
    ap = argparse.ArgumentParser(prog="radix_sort_b200.filebench")
    sub = ap.add_subparsers(dest="cmd", required=True)
    g = sub.add_parser("gen")
    g.add_argument("size", type=float, help="GB of key bytes")
    g.add_argument("file")
    g.add_argument("kind", nargs="?", default="u32")
    b = sub.add_parser("bench")
    b.add_argument("kind")
    b.add_argument("files", nargs="+")
    a = ap.parse_args(argv)
    k = Kind[a.kind.upper()]
    if a.cmd == "gen":
        print(f"{gen_data(a.size, a.file, k)} {a.kind} keys -> {a.file}")
    else:
        secs = bench_sorts_data(a.files, k)
        n = sum(os.path.getsize(f) for f in a.files) // NUMBER_OF_DIGITS[k] / len(a.files)
        print(f"radix_sort: {secs[0]:.6f} s per file ({n / secs[0] / 1e9:.3f} Gkeys/s, host slice in, host slice out)")
    return 0


if __name__ == "__main__":
    raise SystemExit(_main())
