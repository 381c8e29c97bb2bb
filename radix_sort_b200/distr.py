"""Synthetic inputs: the distribution families of the reference's src/distr.rs,
restated as counter-based generators (key i = f(seed, i)).

    KeyUniform / Standard        distr.rs:40-52   -> Distr.UNIFORM
    MyExp(lambda) = Geometric    distr.rs:4-38    -> Distr.GEOMETRIC
    ZipfU32(s) / ZipfU64(s)      distr.rs:54-76,108-130 -> Distr.ZIPF
    StepUniformU32(m) / U64(m)   distr.rs:78-106,132-160 -> Distr.STEP
    presorted                    not in distr.rs (BASELINE.json config 4 asks for
                                 it): ASCENDING / DESCENDING are the builder's
    Uniform[-1,1) + specials     tests.rs:135-143 -> Distr.FLOAT_TEST

`generate` fills a CUDA tensor with the device kernels (K5).  `uniform_host`
is the bit-identical numpy mirror of the UNIFORM stream (integer arithmetic
only), used where a CPU-side input is needed without a GPU.  The reference
seeds nothing (main.rs:27, tests.rs:28): only the families match, not streams.
"""
from __future__ import annotations

import ctypes
import enum

import numpy as np

from . import _ffi
from ._ffi import check
from .api import Kind, NUMBER_OF_DIGITS


class Distr(enum.IntEnum):
    UNIFORM = 0
    GEOMETRIC = 1
    ZIPF = 2
    STEP = 3
    ASCENDING = 4
    DESCENDING = 5
    FLOAT_TEST = 6


_TORCH_DTYPE = {
    Kind.U8: "uint8", Kind.U16: "uint16", Kind.U32: "uint32", Kind.U64: "uint64",
    Kind.I8: "int8", Kind.I16: "int16", Kind.I32: "int32", Kind.I64: "int64",
    Kind.F32: "float32", Kind.F64: "float64", Kind.U128: "uint64", Kind.I128: "uint64",
}


def empty_keys(n: int, kind: Kind, device="cuda"):
    """Uninitialised CUDA tensor for n keys of `kind` ((n, 2) uint64 for 128-bit)."""
    import torch
    dt = getattr(torch, _TORCH_DTYPE[Kind(kind)])
    shape = (n, 2) if NUMBER_OF_DIGITS[Kind(kind)] == 16 else (n,)
    return torch.empty(shape, dtype=dt, device=device)


def generate(out, kind, distr=Distr.UNIFORM, param: float = 0.0, seed: int = 0xB2000001,
             index_base: int = 0) -> None:
    """Fill the CUDA tensor `out` with keys index_base.. of stream (seed, distr, param)."""
    import torch
    k = Kind(kind)
    n = out.numel() * out.element_size() // NUMBER_OF_DIGITS[k]
    with torch.cuda.device(out.device):
        stream = torch.cuda.current_stream(out.device).cuda_stream
        check(_ffi.lib().b200sort_generate(out.data_ptr(), n, int(k), int(distr), float(param),
                                           seed & (2**64 - 1), index_base, ctypes.c_void_p(stream)))


_M1 = np.uint64(0x9E3779B97F4A7C15)
_M2 = np.uint64(0xBF58476D1CE4E5B9)
_M3 = np.uint64(0x94D049BB133111EB)


def mix64(z: np.ndarray) -> np.ndarray:
    """splitmix64 finaliser on a uint64 array (wrap-around arithmetic)."""
    with np.errstate(over="ignore"):
        z = z + _M1
        z = (z ^ (z >> np.uint64(30))) * _M2
        z = (z ^ (z >> np.uint64(27))) * _M3
        return z ^ (z >> np.uint64(31))


def uniform_host(n: int, kind, seed: int = 0xB2000001, index_base: int = 0) -> np.ndarray:
    """numpy mirror of the UNIFORM stream: u = mix64(seed ^ mix64(i)), key = low
    bytes of u (128-bit: lo = u, hi = mix64(u)).  Bit-identical to the device."""
    k = Kind(kind)
    i = np.arange(index_base, index_base + n, dtype=np.uint64)
    u = mix64(np.uint64(seed & (2**64 - 1)) ^ mix64(i))
    kb = NUMBER_OF_DIGITS[k]
    if kb == 16:
        out = np.empty((n, 2), dtype=np.uint64)
        out[:, 0] = u
        out[:, 1] = mix64(u)
        return out
    raw = u.astype({1: np.uint8, 2: np.uint16, 4: np.uint32, 8: np.uint64}[kb])
    return raw.view(np.dtype(_TORCH_DTYPE[k]))
