"""Host-side mirror of the reference's sort interface, over the C ABI.

Reference interface (src/radix_sort/mod.rs):
    pub trait RadixSort<T: RadixDigits> { fn radix_sort(&mut self); }      :18-20
    impl<T: RadixDigits> RadixSort<T> for [T]                               :61
    pub trait RadixSortCopyOnly<T> { fn radix_sort0..5(&mut self); }        :22-38
    pub fn counting_sort(data: &mut [u8])                                   :40-59
Same names, same meaning: in place, blocking, ascending in get_digit's order,
failure = exception (the reference panics, mod.rs:68,106).

`data` is either a host buffer (numpy array: copied to the GPU, sorted, copied
back -- b200sort_keys_host) or a CUDA torch tensor (sorted where it lies --
b200sort_keys_device on the current torch stream).  Every path ends in the
sm_100a kernels of libb200sort.so; nothing here sorts on the CPU.
"""
from __future__ import annotations

import ctypes
import enum
from typing import Optional

import numpy as np

from . import _ffi
from ._ffi import B200SortError, check


class Kind(enum.IntEnum):
    """b200sort_kind: the built-in RadixDigits impls (radix_digits.rs:7-124)."""
    U8 = 0
    U16 = 1
    U32 = 2
    U64 = 3
    U128 = 4
    I8 = 5
    I16 = 6
    I32 = 7
    I64 = 8
    I128 = 9
    F32 = 10
    F64 = 11


USIZE = Kind.U64  # radix_digits.rs:47-53 on a 64-bit target
ISIZE = Kind.I64  # radix_digits.rs:95-101

FLAG_NO_SKIP = 1

# RadixDigits::NUMBER_OF_DIGITS (radix_digits.rs:8,16,24,32,40,56,...,117)
NUMBER_OF_DIGITS = {
    Kind.U8: 1, Kind.U16: 2, Kind.U32: 4, Kind.U64: 8, Kind.U128: 16,
    Kind.I8: 1, Kind.I16: 2, Kind.I32: 4, Kind.I64: 8, Kind.I128: 16,
    Kind.F32: 4, Kind.F64: 8,
}

_NP_KINDS = {
    "uint8": Kind.U8, "uint16": Kind.U16, "uint32": Kind.U32, "uint64": Kind.U64,
    "int8": Kind.I8, "int16": Kind.I16, "int32": Kind.I32, "int64": Kind.I64,
    "float32": Kind.F32, "float64": Kind.F64,
}


def kind_of(data, kind=None) -> Kind:
    """Key kind of a numpy array / torch tensor.  128-bit keys have no numpy or
    torch dtype: pass kind=Kind.U128 / Kind.I128 with an (n, 2) uint64 buffer
    (little-endian: [:, 0] low word, [:, 1] high word)."""
    if kind is not None:
        return Kind(int(kind)) if not isinstance(kind, str) else Kind[kind.upper()]
    if not hasattr(data, "dtype"):
        raise B200SortError(1, f"cannot sort a {type(data).__name__}: pass a numpy array or a CUDA torch tensor")
    name = str(data.dtype).replace("torch.", "")
    try:
        return _NP_KINDS[name]
    except KeyError:
        # a user-defined RadixDigits impl cannot run on the GPU (include/b200sort.h)
        raise B200SortError(2, f"unsupported key type {data.dtype!r}") from None


def key_bytes(kind) -> int:
    return int(_ffi.lib().b200sort_key_bytes(int(kind)))


def temp_bytes(n: int, kind) -> int:
    """Device workspace b200sort_keys_device needs (the ping-pong buffer of
    mod.rs:71-83 plus histograms / bucket bases / look-back state)."""
    return int(_ffi.lib().b200sort_temp_bytes(int(n), int(kind)))


def _is_torch(x) -> bool:
    return type(x).__module__.startswith("torch")


def _count(nbytes: int, kind: Kind) -> int:
    kb = NUMBER_OF_DIGITS[kind]
    if nbytes % kb:
        raise B200SortError(1, "buffer size is not a multiple of the key size")
    return nbytes // kb


def radix_sort(data, kind=None, *, workspace=None, flags: int = 0) -> None:
    """`data.radix_sort()` (mod.rs:61-176): sort `data` in place.

    numpy array -> host drop-in path; CUDA torch tensor -> device path
    (asynchronous on torch's current stream; pass `workspace`, a uint8 CUDA
    tensor of >= temp_bytes(n, kind), to avoid an allocation per call).
    """
    L = _ffi.lib()
    k = kind_of(data, kind)
    if _is_torch(data):
        import torch
        if not data.is_cuda:
            raise B200SortError(1, "torch tensors must live on a CUDA device (use a numpy array for host data)")
        if not data.is_contiguous():
            raise B200SortError(1, "tensor must be contiguous")
        n = _count(data.numel() * data.element_size(), k)
        if n == 0:
            return
        with torch.cuda.device(data.device):
            need = temp_bytes(n, k)
            if workspace is None:
                workspace = torch.empty(need, dtype=torch.uint8, device=data.device)
            elif workspace.numel() * workspace.element_size() < need or workspace.device != data.device:
                raise B200SortError(1, "workspace too small or on another device")
            stream = torch.cuda.current_stream(data.device).cuda_stream
            check(L.b200sort_keys_device_ex(data.data_ptr(), workspace.data_ptr(),
                                            workspace.numel() * workspace.element_size(), n, int(k),
                                            ctypes.c_void_p(stream), flags))
            # the caching allocator may hand the workspace to another stream once
            # freed; tie its lifetime to this stream
            workspace.record_stream(torch.cuda.current_stream(data.device))
        return
    arr = data
    if not isinstance(arr, np.ndarray):
        raise B200SortError(1, f"cannot sort a {type(data).__name__}: pass a numpy array or a CUDA torch tensor")
    if not (arr.flags["C_CONTIGUOUS"] and arr.flags["WRITEABLE"]):
        raise B200SortError(1, "array must be C-contiguous and writable")
    n = _count(arr.size * arr.itemsize, k)
    check(L.b200sort_keys_host(arr.ctypes.data, n, int(k)))


def radix_sort_records(recs, key_kind, key_off: int = 0, *, workspace=None) -> None:
    """`Vec<(K, V)>::radix_sort()` (radix_digits.rs:126-136; tests.rs:175-187,
    main.rs:108-125): stable sort of whole records by the key field.

    recs: numpy structured / (n, 2) array or CUDA tensor of shape (n, 2) whose
    row is one record; key_off = byte offset of the key inside the record.
    """
    L = _ffi.lib()
    k = kind_of(recs, key_kind)
    kb = NUMBER_OF_DIGITS[k]
    rec_bytes = 2 * kb
    if _is_torch(recs):
        import torch
        if not (recs.is_cuda and recs.is_contiguous()):
            raise B200SortError(1, "records tensor must be a contiguous CUDA tensor")
        nbytes = recs.numel() * recs.element_size()
        if nbytes % rec_bytes:
            raise B200SortError(1, "buffer size is not a multiple of the record size")
        n = nbytes // rec_bytes
        if n == 0:
            return
        with torch.cuda.device(recs.device):
            need = int(L.b200sort_records_temp_bytes(n, rec_bytes, int(k)))
            if workspace is None:
                workspace = torch.empty(need, dtype=torch.uint8, device=recs.device)
            stream = torch.cuda.current_stream(recs.device).cuda_stream
            check(L.b200sort_records_device(recs.data_ptr(), workspace.data_ptr(),
                                            workspace.numel() * workspace.element_size(), n, rec_bytes,
                                            key_off, int(k), ctypes.c_void_p(stream)))
            workspace.record_stream(torch.cuda.current_stream(recs.device))
        return
    arr = recs
    if not (isinstance(arr, np.ndarray) and arr.flags["C_CONTIGUOUS"] and arr.flags["WRITEABLE"]):
        raise B200SortError(1, "records must be a C-contiguous writable numpy array")
    nbytes = arr.size * arr.itemsize
    if nbytes % rec_bytes:
        raise B200SortError(1, "buffer size is not a multiple of the record size")
    check(L.b200sort_records_host(arr.ctypes.data, nbytes // rec_bytes, rec_bytes, key_off, int(k)))


# RadixSortCopyOnly (mod.rs:22-38): the reference's experimental single-thread /
# thread-per-digit / native-thread / rayon / buffered variants.  They differ in
# how the CPU work is scheduled, not in the result, so on the GPU they are all
# the same sort.  counting_sort (mod.rs:40-59) is the u8 case: one digit pass.
radix_sort0 = radix_sort1 = radix_sort2 = radix_sort3 = radix_sort4 = radix_sort5 = radix_sort


def counting_sort(data) -> None:
    if kind_of(data) != Kind.U8:
        raise B200SortError(2, "counting_sort takes u8 keys (mod.rs:40)")
    radix_sort(data)


def launch_count() -> int:
    """Kernels launched by the library in this process (bench.py: gpu_launches)."""
    return int(_ffi.lib().b200sort_launch_count())


def set_variant(key_bytes_: int, variant: int) -> None:
    """Select K3 variant `variant` (-1 = default) for keys of key_bytes_ bytes
    (108 / 116: 8- / 16-byte records).  Tuning sweeps only."""
    check(_ffi.lib().b200sort_set_variant(key_bytes_, variant))


def set_host_pipeline_min_bytes(nbytes: int) -> None:
    """Host slices of at least `nbytes` take the pipelined host path (range split
    + copy-out overlapped with the range sorts); 0 = always, 2**64-1 = never."""
    check(_ffi.lib().b200sort_set_host_pipeline_min_bytes(int(nbytes)))


def release_cached_memory() -> None:
    """Free the device blocks the host entry points keep between calls (b200sort_release_cached_memory)."""
    check(_ffi.lib().b200sort_release_cached_memory())


def debug_force_rerank(on: bool) -> None:
    """Test hook: every K3 tile of every later sort takes the ballot re-rank slow path."""
    _ffi.lib().b200sort_debug_force_rerank(1 if on else 0)


def debug_small_sort(on: bool) -> None:
    """Test hook: False sends slices of <= 4096 elements through the big K1-K2-K3 path too."""
    _ffi.lib().b200sort_debug_small_sort(1 if on else 0)


def debug_narrow_fill(on: bool) -> None:
    """Test hook: False sends key-only 1- and 2-byte keys through the onesweep passes instead of
    the histogram + fill path (csrc/narrow_kernels.cuh)."""
    _ffi.lib().b200sort_debug_narrow_fill(1 if on else 0)


def debug_reranked_tiles(workspace, n: int, kind, rec_bytes: int = 0) -> int:
    """Tiles of the last sort run on `workspace` (a CUDA tensor) that took the slow path."""
    import torch
    out = ctypes.c_uint(0)
    stream = torch.cuda.current_stream(workspace.device).cuda_stream
    with torch.cuda.device(workspace.device):
        check(_ffi.lib().b200sort_debug_reranked_tiles(workspace.data_ptr(), workspace.numel(), n, int(Kind(kind)), rec_bytes,
                                                       ctypes.c_void_p(stream), ctypes.byref(out)))
    return int(out.value)


def num_variants(key_bytes_: int) -> int:
    return int(_ffi.lib().b200sort_num_variants(key_bytes_))


def sort_timed(data, kind=None, *, workspace=None, flags: int = 0):
    """Device sort with per-phase CUDA-event timing (synchronises).  Returns a
    dict: hist_ms, scan_ms, pass_ms[D], copy_back_ms, total_ms."""
    import torch
    L = _ffi.lib()
    k = kind_of(data, kind)
    n = _count(data.numel() * data.element_size(), k)
    D = NUMBER_OF_DIGITS[k]
    with torch.cuda.device(data.device):
        need = temp_bytes(n, k)
        if workspace is None:
            workspace = torch.empty(need, dtype=torch.uint8, device=data.device)
        out = (ctypes.c_float * (D + 4))()
        stream = torch.cuda.current_stream(data.device).cuda_stream
        check(L.b200sort_keys_device_timed(data.data_ptr(), workspace.data_ptr(),
                                           workspace.numel() * workspace.element_size(), n, int(k),
                                           ctypes.c_void_p(stream), flags, out, D + 4))
    v = list(out)
    return {"hist_ms": v[0], "scan_ms": v[1], "pass_ms": v[2:2 + D], "copy_back_ms": v[2 + D],
            "total_ms": v[3 + D]}
