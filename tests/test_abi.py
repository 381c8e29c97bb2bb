"""CPU: the C-ABI library builds, loads, and exports every symbol that
include/b200sort.h declares; host-side argument checking; loud failure when no
CUDA device exists (there is no CPU fallback).  No compute calls."""
import ctypes
import os
import re

import numpy as np
import pytest

import radix_sort_b200 as rs
from radix_sort_b200 import _build, _ffi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200sort.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200sort_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_all_exported_and_bound():
    syms = _declared_symbols()
    assert len(syms) >= 15
    L = ctypes.CDLL(_build.build())
    for s in syms:
        assert hasattr(L, s), f"{s} declared in b200sort.h but not exported"
        assert s in _ffi.PROTOTYPES, f"{s} has no ctypes prototype"
    assert sorted(_ffi.PROTOTYPES) == syms


def test_version_and_key_bytes():
    L = _ffi.lib()
    assert L.b200sort_version() == 100
    want = {rs.Kind.U8: 1, rs.Kind.U16: 2, rs.Kind.U32: 4, rs.Kind.U64: 8, rs.Kind.U128: 16,
            rs.Kind.I8: 1, rs.Kind.I16: 2, rs.Kind.I32: 4, rs.Kind.I64: 8, rs.Kind.I128: 16,
            rs.Kind.F32: 4, rs.Kind.F64: 8}
    for k, b in want.items():
        assert rs.key_bytes(k) == b == rs.NUMBER_OF_DIGITS[k]  # NUMBER_OF_DIGITS = byte width
    assert L.b200sort_key_bytes(99) == 0
    assert rs.USIZE == rs.Kind.U64 and rs.ISIZE == rs.Kind.I64


def test_temp_bytes_covers_ping_pong_buffer():
    for k in rs.Kind:
        prev = 0
        for n in (0, 1, 1000, 10**6, 10**9, 3 * 10**9):
            t = rs.temp_bytes(n, k)
            assert t >= n * rs.key_bytes(k) and t >= prev and t % 256 == 0
            prev = t
    assert _ffi.lib().b200sort_temp_bytes(1000, 99) == 0


def test_kind_inference_and_unsupported_types():
    assert rs.kind_of(np.zeros(1, np.uint32)) == rs.Kind.U32
    assert rs.kind_of(np.zeros(1, np.float64)) == rs.Kind.F64
    assert rs.kind_of(np.zeros((1, 2), np.uint64), "i128") == rs.Kind.I128
    with pytest.raises(rs.B200SortError) as e:
        rs.radix_sort(np.zeros(4, np.complex64))
    assert e.value.code == 2  # B200SORT_ERR_UNSUPPORTED_KIND
    with pytest.raises(rs.B200SortError):
        rs.radix_sort([3, 1, 2])
    with pytest.raises(rs.B200SortError):
        rs.radix_sort(np.zeros((4, 4), np.uint32)[:, ::2])


def test_unknown_kind_and_bad_record_shape_rejected_before_cuda():
    L = _ffi.lib()
    buf = np.zeros(16, np.uint8)
    assert L.b200sort_keys_host(buf.ctypes.data, 4, 99) == 2
    assert b"kind" in L.b200sort_last_error()
    assert L.b200sort_records_host(buf.ctypes.data, 2, 12, 0, int(rs.Kind.U32)) == 2
    assert L.b200sort_set_variant(3, 0) == 1
    assert L.b200sort_set_variant(4, 999) == 1
    assert L.b200sort_set_variant(4, -1) == 0


def test_empty_input_is_a_noop_everywhere():
    # the reference panics on n == 0 (mod.rs:92 chunks(0)); documented deviation: no-op
    rs.radix_sort(np.zeros(0, np.uint64))
    rs.radix_sort_records(np.zeros((0, 2), np.uint32), rs.Kind.U32)


def test_no_cpu_fallback_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    a = np.array([3, 1, 2], dtype=np.uint32)
    with pytest.raises(rs.B200SortError) as e:
        rs.radix_sort(a)
    assert e.value.code in (3, 4)
    assert a.tolist() == [3, 1, 2]  # untouched: nothing sorted it on the CPU


def test_communicator_fails_loudly_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    from radix_sort_b200 import mgpu
    with pytest.raises(rs.B200SortError) as e:
        mgpu.Communicator([0])
    assert e.value.code == 3
    with pytest.raises(rs.B200SortError) as e:  # argument errors are answered before any CUDA call
        mgpu.Communicator([0] * 17)
    assert e.value.code == 1


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "radix_sort_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                for pat in (r"^\s*(from|import)\s+oracle", r"liboracle", r"oracle[/\\.]", r"cpu_oracle"):
                    assert not re.search(pat, text, flags=re.M), f"{f} reaches into oracle/ ({pat})"


_C_CALLER = r"""
#include <stdio.h>
#include <string.h>
#include "b200sort.h"

/* what a C (or cgo / Rust `extern "C"`) caller sees: plain pointers and sizes.
 * Only calls that are answered before any CUDA work. */
int main(void) {
    b200sort_comm *comm = (b200sort_comm *)0;
    if (b200sort_version() != B200SORT_VERSION) return 1;
    if (b200sort_key_bytes(B200SORT_U64) != 8 || b200sort_key_bytes(B200SORT_I128) != 16) return 2;
    if (b200sort_key_bytes(B200SORT_KIND_COUNT) != 0) return 3;
    if (b200sort_temp_bytes(1000, B200SORT_F32) < 4000) return 4;
    if (b200sort_keys_host((void *)0, 0, B200SORT_U32) != B200SORT_OK) return 5;          /* empty slice */
    if (b200sort_keys_host((void *)0, 0, 99) != B200SORT_ERR_UNSUPPORTED_KIND) return 6;
    if (strlen(b200sort_last_error()) == 0) return 7;
    if (b200sort_mgpu_create(&comm, 0, (const int *)0) != B200SORT_ERR_INVALID_ARG || comm) return 8;
    if (b200sort_mgpu_create(&comm, 17, (const int *)0) != B200SORT_ERR_INVALID_ARG || comm) return 9;
    if (b200sort_mgpu_keys_host((b200sort_comm *)0, (void *)0, 0, B200SORT_U8) != B200SORT_ERR_INVALID_ARG) return 10;
    if (b200sort_mgpu_destroy((b200sort_comm *)0) != B200SORT_OK) return 11;
    puts("ok");
    return 0;
}
"""


def test_header_is_plain_c_and_a_c_program_links(tmp_path):
    """include/b200sort.h compiles as C99 with -pedantic -Werror, and a C program
    linked against libb200sort.so gets the documented answers (no GPU needed)."""
    import shutil
    import subprocess
    if shutil.which("gcc") is None:
        pytest.skip("no gcc")
    lib = _build.build()
    src = tmp_path / "caller.c"
    src.write_text(_C_CALLER)
    exe = tmp_path / "caller"
    inc = os.path.join(ROOT, "include")
    subprocess.run(["gcc", "-std=c99", "-pedantic", "-Wall", "-Wextra", "-Werror", "-fsyntax-only", "-x", "c",
                    os.path.join(inc, "b200sort.h")], check=True)
    libdir = os.path.dirname(lib)
    subprocess.run(["gcc", "-std=c99", "-pedantic", "-Wall", "-Wextra", "-Werror", "-I", inc, str(src), "-o", str(exe),
                    "-L", libdir, "-lb200sort", f"-Wl,-rpath,{libdir}"], check=True)
    out = subprocess.run([str(exe)], capture_output=True, text=True)
    assert out.returncode == 0, f"check {out.returncode} failed: {out.stderr}"
    assert out.stdout.strip() == "ok"
