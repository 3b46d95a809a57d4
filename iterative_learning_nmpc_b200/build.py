"""Build libnmpc_b200.so (hand-written sm_100a CUDA + C ABI) in-tree with nvcc.

Plays the role of acados' code generation + `cc` step behind `MPCOptConfig.recompile`
(reference: `mpc_controller/utils/solver.py:68-72`, README.md:12).  `regenerate=True` re-runs the
model code generator first (needed only when the URDF changes; the generated header is
committed).
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "libnmpc_b200.so")
SOURCES = ["nmpc_abi.cu"]
DEPS = sorted(f for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh", ".h"))) + \
    [os.path.join("..", "..", "include", "nmpc_b200.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "--shared", "-Xcompiler", "-fPIC"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; libnmpc_b200.so cannot be built")


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(os.path.join(CSRC, d)) > t for d in DEPS)


def build(force: bool = False, regenerate: bool = False, verbose: bool = False) -> str:
    if regenerate:
        from .codegen import emit
        argv = sys.argv
        sys.argv = ["emit"]
        try:
            emit.main()
        finally:
            sys.argv = argv
    if force or needs_build():
        cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB] + SOURCES
        subprocess.check_call(cmd, cwd=CSRC)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, regenerate="--regenerate" in sys.argv,
                verbose="--verbose" in sys.argv))
