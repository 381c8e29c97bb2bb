"""In-tree build of libb200sort.so with nvcc for sm_100a (and nothing else).

The .so is written next to this file (radix_sort_b200/libb200sort.so); it is
git-ignored but travels to the GPU box with the gpurun snapshot.
"""
from __future__ import annotations

import hashlib
import os
import re
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "libb200sort.so")
STAMP = os.path.join(PKG, ".libb200sort.stamp")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
]
FAMILIES = {"KEYS": 0, "RECORDS": 1, "PARTITION": 2}


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libb200sort.so cannot be built (there is no CPU fallback)")


def _fingerprint() -> str:
    h = hashlib.sha256()
    files = sorted(os.listdir(CSRC)) + [os.path.join(ROOT, "include", "b200sort.h")]
    for f in files:
        path = f if os.path.isabs(f) else os.path.join(CSRC, f)
        if os.path.isfile(path):
            h.update(os.path.basename(path).encode())  # names, not absolute paths: the tree moves (gpurun)
            with open(path, "rb") as fh:
                h.update(fh.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def needs_build() -> bool:
    if not os.path.exists(LIB) or not os.path.exists(STAMP):
        return True
    with open(STAMP) as fh:
        return fh.read().strip() != _fingerprint()


def translation_units():
    """(object name, source, extra -D flags) of every translation unit.  The K3
    variants are compiled one kernel per unit (csrc/sort_impl.cuh names how many
    each (family, width) table holds) so that the build parallelises."""
    units = [("api", "b200sort_api.cu", []), ("aux", "aux_kernels.cu", [])]
    text = open(os.path.join(CSRC, "sort_impl.cuh")).read()
    for fam, w, nv in re.findall(r"#define B200_NV_([A-Z]+)_(\d+) (\d+)", text):
        d = [f"-DB200_FAM={FAMILIES[fam]}", f"-DB200_W={w}"]
        units.append((f"front_{fam.lower()}_{w}", "inst_front.cu", d))
        for v in range(int(nv)):
            units.append((f"k3_{fam.lower()}_{w}_v{v}", "inst_variant.cu", d + [f"-DB200_V={v}"]))
    return units


def build(force: bool = False, verbose: bool = False, jobs: int | None = None) -> str:
    """Compile csrc/ into libb200sort.so (sm_100a only).  Returns its path."""
    if not force and not needs_build():
        return LIB
    nvcc = _nvcc()
    objdir = os.path.join(PKG, "build") if verbose else os.path.join(PKG, f"build.{os.getpid()}")
    os.makedirs(objdir, exist_ok=True)
    jobs = jobs or int(os.environ.get("B200SORT_BUILD_JOBS", "0")) or max(1, len(os.sched_getaffinity(0)))
    units = translation_units()
    # the heavy units (many unrolled items) first, so the tail of the build is short
    units.sort(key=lambda u: (not u[0].startswith("k3_"), u[0]))
    log = []

    def compile_one(unit):
        name, src, defs = unit
        obj = os.path.join(objdir, name + ".o")
        cmd = [nvcc, *NVCC_FLAGS, *defs, "-I", os.path.join(ROOT, "include"), "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        return name, obj, p.returncode, p.stdout

    from concurrent.futures import ThreadPoolExecutor
    objs = []
    with ThreadPoolExecutor(max_workers=jobs) as ex:
        for name, obj, rc, out in ex.map(compile_one, units):
            if verbose:
                log.append(f"==== {name}\n{out}")
            if rc:
                sys.stderr.write(out)
                raise RuntimeError(f"nvcc failed on {name} (exit {rc})")
            objs.append(obj)
    if verbose:
        with open(os.path.join(objdir, "ptxas.log"), "w") as fh:
            fh.write("\n".join(log))
    tmp_lib = LIB + f".tmp{os.getpid()}"
    subprocess.run([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", tmp_lib, *objs], check=True)
    os.replace(tmp_lib, LIB)  # atomic: concurrent ranks never see a half-written library
    if not verbose:
        shutil.rmtree(objdir, ignore_errors=True)
    with open(STAMP, "w") as fh:
        fh.write(_fingerprint())
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
