"""Builds soillib_b200/libsoilb200.so in-tree with nvcc for sm_100a.

The library is the C-ABI declared in include/soilb200.h.  -fmad=false keeps the
kernels' arithmetic "as written" (no FMA contraction) so results reproduce
bit-for-bit against the CPU oracle; -lineinfo keeps ncu's source page usable.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libsoilb200.so")
SOURCES = ["flow.cu", "accumulate.cu", "accumulate_tiled.cu", "transport.cu", "dense.cu", "noise.cu", "multiflow.cu", "grad.cu", "albedo.cu", "path.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++20", "-fmad=false",
    "-Xcompiler", "-fPIC,-O3,-ffp-contract=off", "-shared",
]


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "soilb200.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB
    cmd = [NVCC] + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB] + [os.path.join(CSRC, s) for s in SOURCES]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        sys.stderr.write(proc.stdout + proc.stderr)
        raise RuntimeError("nvcc failed building libsoilb200.so")
    if verbose:
        sys.stderr.write(proc.stderr)
    return LIB


if __name__ == "__main__":
    print(build(force=True, verbose="-v" in sys.argv))
