"""File-based bench input mode (reference main.rs:47-99: gen_data / bench_sorts_data)."""
import os

import numpy as np
import pytest

import radix_sort_b200 as rs
from radix_sort_b200 import distr, filebench

from . import _ref


@pytest.mark.parametrize("kind,size", [("u32", 0.001), ("u64", 0.0005), ("u16", 0.0001), ("u128", 0.0004)])
def test_gen_data_is_a_raw_native_endian_dump(kind, size, tmp_path):
    k = rs.Kind[kind.upper()]
    kb = rs.key_bytes(k)
    path = tmp_path / f"{kind}.bin"
    n = filebench.gen_data(size, path, k, seed=7)
    assert n == int(np.float32(size) * np.float32(1e9) / np.float32(kb))  # main.rs:88, f32 arithmetic
    assert os.path.getsize(path) == n * kb  # no header
    got = filebench.load_data(path, k)
    want = distr.uniform_host(n, k, seed=7)
    assert got.tobytes() == want.tobytes()


def test_load_data_drops_a_trailing_partial_element(tmp_path):
    path = tmp_path / "odd.bin"
    with open(path, "wb") as fh:
        fh.write(bytes(range(23)))  # 5 u32 + 3 stray bytes: data.len() / size_of::<T>() (main.rs:59)
    a = filebench.load_data(path, rs.Kind.U32)
    assert a.shape == (5,) and a.tobytes() == bytes(range(20))


def test_bench_sorts_data_without_files_or_sorts():
    assert filebench.bench_sorts_data([], rs.Kind.U32) == []  # main.rs:51-53
    assert filebench.bench_sorts_data(["x"], rs.Kind.U32, sorts=()) == []


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["u32", "i64", "f32"])
def test_bench_sorts_data_sorts_every_file(kind, tmp_path, cuda):
    k = rs.Kind[kind.upper()]
    files = []
    for j in range(2):
        p = tmp_path / f"{kind}_{j}.bin"
        filebench.gen_data(0.002, p, k, seed=11 + j)
        files.append(p)
    seen = []

    def sort_and_keep(a, kk):
        before = a.copy()
        rs.radix_sort(a, kk)
        seen.append((before, a.copy()))

    secs = filebench.bench_sorts_data(files, k, sorts=(sort_and_keep,))
    assert len(secs) == 1 and secs[0] > 0
    assert len(seen) == 2
    for before, after in seen:
        assert _ref.same_bits(after, _ref.expected_sorted(before, kind))
