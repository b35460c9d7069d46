"""Builds libupow_b200.so in-tree with nvcc for sm_100a (no torch, no JIT cache)."""
from __future__ import annotations

import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
# UPOW_CHECKED=1 selects the build with device-side index asserts (tools/sanitizer_target.py)
CHECKED = os.environ.get("UPOW_CHECKED", "") not in ("", "0")
LIB = os.path.join(HERE, "libupow_b200_checked.so" if CHECKED else "libupow_b200.so")
# experiments: UPOW_LIB=<path> builds / loads another library file, UPOW_NVCC_EXTRA="<flags>" is appended to the nvcc line
if os.environ.get("UPOW_LIB"):
    LIB = os.path.abspath(os.environ["UPOW_LIB"])
NVCC_EXTRA = os.environ.get("UPOW_NVCC_EXTRA", "").split()
SOURCES = ["upow_lib.cu", "microbench.cu", "noalloc.cu"]
HEADERS = ["kernels.cuh", "search_kernels.cuh", "hot_loop.h", "sha256_core.h", "template_prep.h", os.path.join("..", "..", "include", "upow_b200.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
              "-Xcompiler", "-fPIC,-O2,-Wall", "-shared", "-cudart", "static", "--threads", "3"]


def _nvcc() -> str:
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found: libupow_b200.so cannot be built (there is no CPU fallback)")
    return nvcc


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB
    all_variants = os.environ.get("UPOW_ALL_VARIANTS", "") not in ("", "0")   # the 13 slower builds of round 1 too
    cmd = [_nvcc()] + NVCC_FLAGS + (["-DUPOW_CHECKED"] if CHECKED else []) + (["-DUPOW_ALL_VARIANTS"] if all_variants else []) + (["-Xptxas", "-v"] if verbose else []) + NVCC_EXTRA + \
          ["-I", CSRC, "-o", LIB] + [os.path.join(CSRC, s) for s in SOURCES]
    subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    import sys
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
