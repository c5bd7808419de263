"""Builds libtfpstore.so (CUDA, sm_100a) in-tree with nvcc.

The library is the product: there is no other implementation behind the Python
or Go hosts, so a failed build is a hard error.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libtfpstore.so")
SOURCES = [os.path.join(CSRC, "tfp_store.cu")]
HEADERS = [os.path.join(CSRC, f) for f in ("common.cuh", "select.cuh", "scan.cuh", "mma_scan.cuh", "ingest.cuh", "host_index.hpp", "multi_store.inl")] + [
    os.path.join(os.path.dirname(HERE), "include", "tfp_store.h")
]


def nvcc_path() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; libtfpstore.so cannot be built")


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(f) > t for f in SOURCES + HEADERS + [os.path.abspath(__file__)])


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB
    # Several processes may get here at once (the ranks of a multi-GPU run on a box whose copy of the library is
    # older than a source file): one builds, the others wait for it and then find the library up to date.
    import fcntl
    with open(os.path.join(HERE, ".build.lock"), "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        if not force and not needs_build():
            return LIB
        return _build_locked(verbose)


def _build_locked(verbose: bool) -> str:
    tmp = LIB + ".tmp.%d" % os.getpid()
    cmd = [
        nvcc_path(), "-shared", "-o", tmp, *SOURCES,
        "-gencode", "arch=compute_100a,code=sm_100a",
        "-O3", "-std=c++17", "-lineinfo",
        "-Xcompiler", "-fPIC,-O3,-ffp-contract=off,-Wall,-Wno-unknown-pragmas",
        "--fmad=false",          # scores must round like the reference: no FMA contraction anywhere
        "-Xptxas", "-v" if verbose else "-O3",
        "-cudart", "static",
    ]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
    if res.returncode != 0:
        if os.path.exists(tmp):
            os.remove(tmp)
        raise RuntimeError("nvcc failed building libtfpstore.so")
    os.replace(tmp, LIB)      # a process that has the old library mapped keeps its inode; nobody sees a half-written file
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
