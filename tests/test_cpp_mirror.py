"""include/radix_sort.hpp -- the compiled-language mirror of `slice.radix_sort()`
-- and tests/cpp/tests_rs.cpp, the reference's 15 tests (tests.rs) restated over it.

CPU: the harness compiles warning-free, its logic is checked against the oracle
(tests/cpp/oracle_shim.c answers the three C-ABI symbols the mirror uses), it
notices a sort that does not sort, and it links against libb200sort.so.
GPU: the same program linked against libb200sort.so, run as is."""
import os
import shutil
import subprocess

import pytest

from oracle import cpu_oracle
from radix_sort_b200 import _build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CPP = os.path.join(ROOT, "tests", "cpp")
INC = os.path.join(ROOT, "include")
WARN = ["-O2", "-Wall", "-Wextra", "-Werror", "-pthread"]

pytestmark = pytest.mark.skipif(shutil.which("g++") is None, reason="no g++")


def _compile_harness(out_dir):
    obj = os.path.join(out_dir, "tests_rs.o")
    subprocess.run(["g++", "-std=c++17", *WARN, "-I", INC, "-c", os.path.join(CPP, "tests_rs.cpp"), "-o", obj], check=True)
    return obj


def _link(obj, extra, exe, libdir, lib):
    subprocess.run(["g++", obj, *extra, "-o", exe, "-L", libdir, f"-l{lib}", f"-Wl,-rpath,{libdir}", "-pthread"], check=True)
    return exe


def _run(exe, *args):
    return subprocess.run([exe, *args], capture_output=True, text=True, timeout=600)


def test_harness_logic_against_the_oracle(tmp_path):
    cpu_oracle.build()
    obj = _compile_harness(str(tmp_path))
    shim = str(tmp_path / "oracle_shim.o")
    subprocess.run(["gcc", "-std=c99", *WARN, "-c", os.path.join(CPP, "oracle_shim.c"), "-o", shim], check=True)
    exe = _link(obj, [shim], str(tmp_path / "tests_rs_oracle"), os.path.join(ROOT, "oracle"), "oracle")
    for args in ((), ("--test-threads=1",)):
        out = _run(exe, *args)
        assert out.returncode == 0, out.stdout + out.stderr
        assert out.stdout.count(" ... ok") == 15
        assert "test result: ok. 15 passed; 0 failed" in out.stdout


def test_harness_notices_a_sort_that_does_not_sort(tmp_path):
    obj = _compile_harness(str(tmp_path))
    lazy = tmp_path / "lazy.c"
    lazy.write_text(
        "#include <stddef.h>\n"
        "int b200sort_keys_host(void *k, size_t n, int kind) { (void)k; (void)n; (void)kind; return 0; }\n"
        "int b200sort_records_host(void *r, size_t n, size_t b, size_t o, int kind) { (void)r; (void)n; (void)b; (void)o; (void)kind; return 3; }\n"
        "const char *b200sort_last_error(void) { return \"lazy\"; }\n")
    lazy_o = str(tmp_path / "lazy.o")
    subprocess.run(["gcc", "-std=c99", *WARN, "-c", str(lazy), "-o", lazy_o], check=True)
    exe = str(tmp_path / "tests_rs_lazy")
    subprocess.run(["g++", obj, lazy_o, "-o", exe, "-pthread"], check=True)
    out = _run(exe)
    assert out.returncode == 15
    assert out.stdout.count("FAILED (Not sorted properly!)") == 14
    assert "radix_sort_tuple ... FAILED (b200sort error 3: lazy)" in out.stdout  # a non-zero status is the panic


def test_mirror_compile_time_facts():
    """NUMBER_OF_DIGITS and kind of every built-in impl (radix_digits.rs:7-136); types
    without a get_digit have no RadixDigits (static_asserts: compiling is the test)."""
    subprocess.run(["g++", "-std=c++17", "-Wall", "-Wextra", "-Werror", "-I", INC, "-fsyntax-only",
                    os.path.join(CPP, "mirror_static.cpp")], check=True)


def test_mirror_links_against_the_library(tmp_path):
    lib = _build.build()
    obj = _compile_harness(str(tmp_path))
    _link(obj, [], str(tmp_path / "tests_rs"), os.path.dirname(lib), "b200sort")


@pytest.mark.gpu
def test_reference_test_suite_in_cpp_on_the_gpu(tmp_path):
    """tests.rs, all 15, through radix_sort.hpp -> libb200sort.so on the GPU, on
    parallel threads as libtest runs them, then once more serially."""
    lib = _build.build()
    obj = _compile_harness(str(tmp_path))
    exe = _link(obj, [], str(tmp_path / "tests_rs"), os.path.dirname(lib), "b200sort")
    for args in ((), ("--test-threads=1",)):
        out = _run(exe, *args)
        assert out.returncode == 0, out.stdout + out.stderr
        assert "test result: ok. 15 passed; 0 failed" in out.stdout
