"""The Rust shim (shim/) cannot be compiled here (no rustc: SURVEY.md section 0, fact 10), so this
checks what CAN be checked without a Rust toolchain: every `extern "C"` declaration of
shim/src/radix_sort/ffi.rs names a symbol include/b200sort.h declares, with the same number of
arguments and the same pointer / integer class for every argument and the result; the Kind enum
carries the header's values; the library exports every one of those symbols; and mod.rs.patch
still applies to the reference where the reference tree is present (this container only)."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200sort.h")
FFI_RS = os.path.join(ROOT, "shim", "src", "radix_sort", "ffi.rs")


def _strip_comments(text, line_prefix):
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    return "\n".join(l.split(line_prefix)[0] for l in text.splitlines())


def header_prototypes():
    text = _strip_comments(open(HEADER).read(), "//")
    protos = {}
    for m in re.finditer(r"([A-Za-z_][\w\s\*]*?)\b(b200sort_\w+)\s*\(([^;{]*?)\)\s*;", text, flags=re.S):
        ret, name, args = m.group(1).strip(), m.group(2), m.group(3).strip()
        arglist = [] if args in ("", "void") else [a.strip() for a in args.split(",")]
        protos[name] = ("ptr" if "*" in ret else ("void" if ret.split()[-1] == "void" else "int"),
                        ["ptr" if "*" in a else "int" for a in arglist])
    return protos


def rust_prototypes():
    text = _strip_comments(open(FFI_RS).read(), "//")
    block = re.search(r'extern\s+"C"\s*\{(.*?)\n\}', text, flags=re.S).group(1)
    protos = {}
    for m in re.finditer(r"pub\s+fn\s+(b200sort_\w+)\s*\(([^)]*)\)\s*(?:->\s*([^;]+))?;", block, flags=re.S):
        name, args, ret = m.group(1), m.group(2).strip(), (m.group(3) or "").strip()
        arglist = [a.strip() for a in args.split(",") if a.strip()]
        types = [a.split(":", 1)[1].strip() for a in arglist]
        protos[name] = ("void" if not ret else ("ptr" if ret.startswith("*") else "int"),
                        ["ptr" if t.startswith("*") else "int" for t in types])
    return protos


def test_every_rust_declaration_matches_the_header():
    h, r = header_prototypes(), rust_prototypes()
    assert len(r) >= 10, "ffi.rs parse found too few declarations"
    for name, (ret, args) in r.items():
        assert name in h, f"{name}: declared in ffi.rs but not in include/b200sort.h"
        hret, hargs = h[name]
        assert len(args) == len(hargs), f"{name}: {len(args)} arguments in ffi.rs, {len(hargs)} in the header"
        assert args == hargs, f"{name}: argument classes differ: ffi.rs {args}, header {hargs}"
        assert ret == hret, f"{name}: result class differs: ffi.rs {ret}, header {hret}"
    # the entry points the GPU impl of RadixSort::radix_sort calls (shim/src/radix_sort/gpu.rs)
    gpu = open(os.path.join(ROOT, "shim", "src", "radix_sort", "gpu.rs")).read()
    for used in re.findall(r"ffi::(b200sort_\w+)", gpu):
        assert used in r, f"gpu.rs calls {used}, which ffi.rs does not declare"


def test_kind_enum_matches_the_header():
    h = open(HEADER).read()
    enum = re.search(r"typedef enum b200sort_kind\s*\{(.*?)\}", h, flags=re.S).group(1)
    want = {m.group(1): int(m.group(2)) for m in re.finditer(r"B200SORT_(\w+)\s*=\s*(\d+)", enum)}
    want.pop("KIND_COUNT")
    rs = open(FFI_RS).read()
    body = re.search(r"pub enum Kind\s*\{(.*?)\}", rs, flags=re.S).group(1)
    got = {m.group(1): int(m.group(2)) for m in re.finditer(r"(\w+)\s*=\s*(\d+)", body)}
    assert got == want and len(got) == 12


def test_library_exports_every_declared_symbol():
    """(no compute call: symbol lookup only -- runs without a GPU)"""
    import ctypes
    from radix_sort_b200 import _build
    lib = ctypes.CDLL(_build.build())
    for name in rust_prototypes():
        getattr(lib, name)


def test_patch_applies_to_the_reference(tmp_path):
    ref = "/root/reference"
    if not os.path.isdir(ref) or shutil.which("patch") is None:
        pytest.skip("the reference tree exists only in the build container")
    crate = tmp_path / "radix_sort"
    shutil.copytree(ref, crate)
    for d, _, fs in os.walk(crate):
        os.chmod(d, 0o755)
        for f in fs:
            os.chmod(os.path.join(d, f), 0o644)
    p = subprocess.run(["sh", os.path.join(ROOT, "shim", "apply.sh"), str(crate)], capture_output=True, text=True)
    assert p.returncode == 0, p.stdout + p.stderr
    mod = (crate / "src" / "radix_sort" / "mod.rs").read_text()
    assert '#[cfg(feature = "b200")]\nmod gpu;' in mod
    assert '#[cfg(not(feature = "b200"))]\nimpl<T: RadixDigits> RadixSort<T> for [T]' in mod
    assert (crate / "build.rs").exists() and (crate / "src" / "radix_sort" / "ffi.rs").exists()
    # untouched: the reference's tests and call sites
    for f in ("src/radix_sort/tests.rs", "src/main.rs", "src/radix_sort/radix_digits.rs", "src/distr.rs"):
        assert (crate / f).read_bytes() == open(os.path.join(ref, f), "rb").read()


def test_tuple_impl_is_bounded_like_the_reference_and_by_the_library_shapes():
    """radix_digits.rs:126-130 implements RadixDigits for (T, U) where T: RadixDigits + Send + Sync; GpuKey has
    RadixDigits as a supertrait, so its tuple impl needs the same bounds.  And the shapes it admits are exactly
    those check_record_shape (b200sort_api.cu) admits: refused at compile time, not by a panic at run time."""
    gpu = open(os.path.join(ROOT, "shim", "src", "radix_sort", "gpu.rs")).read()
    assert re.search(r"impl<T:\s*GpuKey\s*\+\s*Send\s*\+\s*Sync,\s*U:\s*Send\s*\+\s*Sync>\s*GpuKey\s+for\s+\(T,\s*U\)", gpu)
    ref = open("/root/reference/src/radix_sort/radix_digits.rs").read() if os.path.exists("/root/reference") else None
    if ref is not None:
        assert re.search(r"impl<T, U> RadixDigits for \(T, U\)\s*where\s*T:\s*RadixDigits\s*\+\s*Send\s*\+\s*Sync,\s*U:\s*Send\s*\+\s*Sync,", ref)
    api = open(os.path.join(ROOT, "radix_sort_b200", "csrc", "b200sort_api.cu")).read()
    assert "rec_bytes == 8 && kb == 4" in api and "rec_bytes == 16 && kb == 8" in api
    assert "(rec == 8 && key == 4) || (rec == 16 && key == 8)" in gpu
    hpp = open(os.path.join(ROOT, "include", "radix_sort.hpp")).read()
    assert "(sizeof(R) == 8 && sizeof(K) == 4) || (sizeof(R) == 16 && sizeof(K) == 8)" in hpp
