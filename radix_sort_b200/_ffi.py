"""ctypes binding of libb200sort.so -- one prototype per symbol of include/b200sort.h.

The library is the product: if it cannot be built or loaded this module raises.
There is no CPU fallback anywhere in radix_sort_b200/.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import (POINTER, c_char_p, c_double, c_float, c_int, c_size_t, c_uint,
                    c_uint64, c_ulonglong, c_void_p)

from . import _build

# every symbol include/b200sort.h declares: name -> (restype, argtypes)
PROTOTYPES = {
    "b200sort_version": (c_int, []),
    "b200sort_last_error": (c_char_p, []),
    "b200sort_key_bytes": (c_size_t, [c_int]),
    "b200sort_temp_bytes": (c_size_t, [c_size_t, c_int]),
    "b200sort_keys_device": (c_int, [c_void_p, c_void_p, c_size_t, c_size_t, c_int, c_void_p]),
    "b200sort_keys_device_ex": (c_int, [c_void_p, c_void_p, c_size_t, c_size_t, c_int, c_void_p, c_uint]),
    "b200sort_keys_device_noalloc": (c_int, [c_void_p, c_size_t, c_int, c_void_p]),
    "b200sort_keys_device_pooled": (c_int, [c_void_p, c_size_t, c_int, c_void_p]),
    "b200sort_keys_host": (c_int, [c_void_p, c_size_t, c_int]),
    "b200sort_set_host_pipeline_min_bytes": (c_int, [c_size_t]),
    "b200sort_release_cached_memory": (c_int, []),
    "b200sort_records_temp_bytes": (c_size_t, [c_size_t, c_size_t, c_int]),
    "b200sort_records_device": (c_int, [c_void_p, c_void_p, c_size_t, c_size_t, c_size_t, c_size_t, c_int, c_void_p]),
    "b200sort_records_host": (c_int, [c_void_p, c_size_t, c_size_t, c_size_t, c_int]),
    "b200sort_launch_count": (c_ulonglong, []),
    "b200sort_set_variant": (c_int, [c_int, c_int]),
    "b200sort_num_variants": (c_int, [c_int]),
    "b200sort_keys_device_timed": (c_int, [c_void_p, c_void_p, c_size_t, c_size_t, c_int, c_void_p, c_uint,
                                           POINTER(c_float), c_int]),
    "b200sort_generate": (c_int, [c_void_p, c_size_t, c_int, c_int, c_double, c_uint64, c_uint64, c_void_p]),
    "b200sort_fingerprint": (c_int, [c_void_p, c_size_t, c_int, c_void_p, c_void_p]),
    "b200sort_count_unsorted": (c_int, [c_void_p, c_size_t, c_int, c_void_p, c_void_p]),
    "b200sort_partition_device": (c_int, [c_void_p, c_void_p, c_void_p, c_size_t, c_size_t, c_int, c_void_p, c_int,
                                          c_void_p, c_void_p]),
    "b200sort_partition_count_device": (c_int, [c_void_p, c_void_p, c_size_t, c_size_t, c_int, c_void_p, c_int,
                                                c_void_p, c_void_p]),
    "b200sort_partition_exchange_device": (c_int, [c_void_p, c_void_p, c_size_t, c_size_t, c_int, c_void_p, c_int,
                                                   c_void_p, c_void_p, c_void_p]),
    "b200sort_partition_claim_device": (c_int, [c_void_p, c_void_p, c_size_t, c_size_t, c_int, c_void_p, c_int, c_void_p, c_void_p,
                                                c_void_p, c_void_p, c_void_p]),
    "b200sort_debug_force_rerank": (None, [c_int]),
    "b200sort_debug_small_sort": (None, [c_int]),
    "b200sort_debug_narrow_fill": (None, [c_int]),
    "b200sort_debug_reranked_tiles": (c_int, [c_void_p, c_size_t, c_size_t, c_int, c_size_t, c_void_p, POINTER(c_uint)]),
    "b200sort_partition_exchange_part_device": (c_int, [c_void_p, c_void_p, c_size_t, c_size_t, c_int, c_void_p, c_int,
                                                        c_void_p, c_void_p, ctypes.c_uint, c_int, c_void_p]),
    "b200sort_mgpu_create": (c_int, [POINTER(c_void_p), c_int, POINTER(c_int)]),
    "b200sort_mgpu_keys_device": (c_int, [c_void_p, POINTER(c_void_p), POINTER(c_size_t), c_int, POINTER(c_void_p),
                                          POINTER(c_size_t)]),
    "b200sort_mgpu_destroy": (c_int, [c_void_p]),
    "b200sort_mgpu_keys_host": (c_int, [c_void_p, c_void_p, c_size_t, c_int]),
    "b200sort_mgpu_set_host_min_bytes": (c_int, [c_size_t]),
}

_lib = None


class B200SortError(RuntimeError):
    """Non-zero status from the C ABI.  The Rust shim panics here (the reference
    has no Result: mod.rs:68,106 `expect`); the Python mirror raises."""

    def __init__(self, code: int, message: str):
        super().__init__(f"b200sort error {code}: {message}")
        self.code = code


def lib() -> ctypes.CDLL:
    """Load (building first if the sources changed) libb200sort.so."""
    global _lib
    if _lib is None:
        path = _build.LIB
        if _build.needs_build():
            try:
                path = _build.build()
            except Exception as exc:
                if not os.path.exists(_build.LIB):
                    raise RuntimeError(
                        "libb200sort.so is missing and could not be built; the CUDA library is "
                        "the only implementation (no CPU fallback)") from exc
                # a library older than the sources would run old kernels under new prototypes
                if os.environ.get("B200SORT_ALLOW_STALE") != "1":
                    raise RuntimeError(
                        "the sources of libb200sort.so changed and the rebuild failed; refusing to load the stale "
                        "library (set B200SORT_ALLOW_STALE=1 to load it anyway)") from exc
                import warnings
                warnings.warn("loading a STALE libb200sort.so (B200SORT_ALLOW_STALE=1): the sources changed and the rebuild failed")
        L = ctypes.CDLL(path)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(L, name)  # AttributeError if the .so lacks a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise B200SortError(rc, lib().b200sort_last_error().decode("utf-8", "replace"))
