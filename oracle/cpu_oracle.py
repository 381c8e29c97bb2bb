"""ctypes loader for the CPU oracle (oracle/oracle.c).

TEST INFRASTRUCTURE ONLY -- see the header of oracle.c.  Importable from
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
leg; never from radix_sort_b200/ (the product).

Parity status: "parity unpinned" (no golden vectors in the reference, no Rust
toolchain here); pinned to the property the reference's tests assert, see
tests/test_oracle.py.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")

# kind ids shared with include/b200sort.h
KINDS = {
    "u8": 0, "u16": 1, "u32": 2, "u64": 3, "u128": 4,
    "i8": 5, "i16": 6, "i32": 7, "i64": 8, "i128": 9,
    "f32": 10, "f64": 11,
    "pair_u32": 100, "pair_u64": 101,
}
KEY_BYTES = {
    "u8": 1, "u16": 2, "u32": 4, "u64": 8, "u128": 16,
    "i8": 1, "i16": 2, "i32": 4, "i64": 8, "i128": 16,
    "f32": 4, "f64": 8, "pair_u32": 8, "pair_u64": 16,
}


def build(force: bool = False) -> str:
    """Compile oracle.c with the committed Makefile (gcc only, seconds)."""
    src = os.path.join(_HERE, "oracle.c")
    stale = (not os.path.exists(_SO)) or os.path.getmtime(_SO) < os.path.getmtime(src)
    if force or stale:
        subprocess.run(["make", "-C", _HERE, "-B", "liboracle.so"], check=True,
                       stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_SO)
        L.oracle_sort0.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int]
        L.oracle_sort0.restype = ctypes.c_int
        for fn in (L.oracle_sort_mt, L.oracle_sort4_mt):
            fn.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_int]
            fn.restype = ctypes.c_int
        L.oracle_get_digit.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_uint]
        L.oracle_get_digit.restype = ctypes.c_int
        L.oracle_available_parallelism.restype = ctypes.c_int
        L.oracle_sort0_records.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_int]
        L.oracle_sort0_records.restype = ctypes.c_int
        L.oracle_generate_uniform.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_uint64, ctypes.c_uint64]
        L.oracle_generate_uniform.restype = ctypes.c_int
        _lib = L
    return _lib


def _check(buf: np.ndarray, kind: str) -> int:
    if not (buf.flags["C_CONTIGUOUS"] and buf.flags["WRITEABLE"]):
        raise ValueError("oracle needs a writable C-contiguous buffer")
    nbytes = buf.size * buf.itemsize
    s = KEY_BYTES[kind]
    if nbytes % s:
        raise ValueError("buffer size is not a multiple of the element size")
    return nbytes // s


def sort0(buf: np.ndarray, kind: str) -> None:
    """In-place sequential LSD sort (restates mod.rs:182-212)."""
    n = _check(buf, kind)
    rc = lib().oracle_sort0(buf.ctypes.data, n, KINDS[kind])
    if rc:
        raise RuntimeError(f"oracle_sort0 rc={rc}")


def sort_mt(buf: np.ndarray, kind: str, threads: int | None = None) -> None:
    """In-place threaded production sort (restates mod.rs:61-176)."""
    n = _check(buf, kind)
    rc = lib().oracle_sort_mt(buf.ctypes.data, n, KINDS[kind], threads or available_parallelism())
    if rc:
        raise RuntimeError(f"oracle_sort_mt rc={rc}")


def sort4_mt(buf: np.ndarray, kind: str, threads: int | None = None) -> None:
    """In-place 'rayon' variant (restates mod.rs:405-466)."""
    n = _check(buf, kind)
    rc = lib().oracle_sort4_mt(buf.ctypes.data, n, KINDS[kind], threads or available_parallelism())
    if rc:
        raise RuntimeError(f"oracle_sort4_mt rc={rc}")


def sort0_records(buf: np.ndarray, rec_bytes: int, key_off: int, kind: str) -> None:
    """In-place sequential LSD sort of (T, U) records of any shape: whole records of
    rec_bytes bytes move, the key of kind `kind` sits at byte key_off
    (radix_digits.rs:126-136 over mod.rs:182-212).  Stable."""
    if not (buf.flags["C_CONTIGUOUS"] and buf.flags["WRITEABLE"]):
        raise ValueError("oracle needs a writable C-contiguous buffer")
    nbytes = buf.size * buf.itemsize
    if nbytes % rec_bytes:
        raise ValueError("buffer size is not a multiple of the record size")
    rc = lib().oracle_sort0_records(buf.ctypes.data, nbytes // rec_bytes, rec_bytes, key_off, KINDS[kind])
    if rc:
        raise RuntimeError(f"oracle_sort0_records rc={rc}")


def generate_uniform(n: int, kind: str, seed: int, index_base: int = 0) -> np.ndarray:
    """The UNIFORM input family, generated on the host cores: the same
    stream as radix_sort_b200.distr.uniform_host and the device generator."""
    kb = KEY_BYTES[kind]
    dt = {1: np.uint8, 2: np.uint16, 4: np.uint32, 8: np.uint64, 16: np.uint64}[kb]
    out = np.empty((n, 2) if kb == 16 else (n,), dtype=dt)
    rc = lib().oracle_generate_uniform(out.ctypes.data, n, kb, seed & (2**64 - 1), index_base)
    if rc:
        raise RuntimeError(f"oracle_generate_uniform rc={rc}")
    return out


def get_digit(elem_bytes: bytes, kind: str, index: int) -> int:
    b = ctypes.create_string_buffer(elem_bytes, len(elem_bytes))
    return lib().oracle_get_digit(ctypes.addressof(b), KINDS[kind], index)


def available_parallelism() -> int:
    """available_parallelism() as the reference reads it (mod.rs:66-70): the
    cores this process may run on, not the machine total."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:  # pragma: no cover
        return lib().oracle_available_parallelism()
