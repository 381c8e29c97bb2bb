"""On-device verification (K6): sortedness under get_digit's order and an
order-independent multiset fingerprint.

For key-only sorts the ordered image is a bijection of the bit pattern
(radix_digits.rs:7-124), so "output sorted AND same multiset as the input" is
equivalent to "bit-identical to the reference's output" -- this is how runs too
large for the CPU oracle (1 B .. 16 B keys) are checked.
"""
from __future__ import annotations

import ctypes

from . import _ffi
from ._ffi import check
from .api import Kind, NUMBER_OF_DIGITS


def _n(t, k: Kind) -> int:
    return t.numel() * t.element_size() // NUMBER_OF_DIGITS[k]


def fingerprint(keys, kind):
    """(n, sum mix64(key), xor mix64(key), sum low64(key)) of a CUDA tensor, as ints."""
    import torch
    k = Kind(kind)
    with torch.cuda.device(keys.device):
        out = torch.empty(4, dtype=torch.uint64, device=keys.device)
        stream = torch.cuda.current_stream(keys.device).cuda_stream
        check(_ffi.lib().b200sort_fingerprint(keys.data_ptr(), _n(keys, k), int(k), out.data_ptr(),
                                              ctypes.c_void_p(stream)))
        return tuple(int(x) for x in out.cpu().numpy())


def count_unsorted(keys, kind) -> int:
    """#{i : key[i] > key[i+1]} in get_digit's order; 0 <=> sorted."""
    import torch
    k = Kind(kind)
    with torch.cuda.device(keys.device):
        out = torch.empty(1, dtype=torch.uint64, device=keys.device)
        stream = torch.cuda.current_stream(keys.device).cuda_stream
        check(_ffi.lib().b200sort_count_unsorted(keys.data_ptr(), _n(keys, k), int(k), out.data_ptr(),
                                                 ctypes.c_void_p(stream)))
        return int(out.cpu().numpy()[0])
