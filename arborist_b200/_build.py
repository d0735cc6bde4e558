"""Builds ``csrc/libarborist_b200.so`` (hand-written CUDA for sm_100a, C ABI in
``include/arborist_b200.h``) in-tree with nvcc.  The built library travels with the
repo snapshot; nothing is JIT-compiled at run time."""
from __future__ import annotations

import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(CSRC, "libarborist_b200.so")
SOURCES = ["arb_ingest.cu", "arb_fit.cu", "arb_base.cu"]
HEADERS = ["arb_common.cuh", os.path.join("..", "..", "include", "arborist_b200.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC,-fvisibility=hidden", "-shared", "-cudart", "static"]


def nvcc_path():
    return shutil.which("nvcc") or ("/usr/local/cuda/bin/nvcc"
                                    if os.path.exists("/usr/local/cuda/bin/nvcc") else None)


def is_stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in SOURCES + HEADERS)


def build_library(force: bool = False, verbose: bool = False) -> str:
    if not force and not is_stale():
        return LIB
    nvcc = nvcc_path()
    if nvcc is None:
        raise RuntimeError("nvcc not found: cannot build libarborist_b200.so")
    cmd = [nvcc, *NVCC_FLAGS, *os.environ.get("ARB_NVCC_EXTRA", "").split(), *SOURCES, "-o", LIB]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    res = subprocess.run(cmd, cwd=CSRC, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    return LIB


if __name__ == "__main__":
    import sys

    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv))
