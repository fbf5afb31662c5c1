"""Build the oracle's C restatement (``oracle/csrc/oracle_cpu.c``) and, when the
reference compiles from its own sources, ``oracle/_ref``.  TEST INFRASTRUCTURE ONLY.

``oracle/_ref``: the reference's compiled code for this path is one C++17
translation unit (``vllm_metal/metal/paged_ops.cpp``) that links MLX private
headers, nanobind and Apple frameworks (Metal, IOKit, CoreFoundation;
``vllm_metal/metal/build.py:246-297``) plus Metal shaders that need ``xcrun
metal``.  None of that exists on Linux: the reference is **unbuildable here** and
no ``oracle/_ref`` artefact is produced (stated in DESIGN.md).  The CPU baseline
is therefore this C port (``cpu_baseline.kind = "port"``).
"""

from __future__ import annotations

import ctypes
import os
import shutil
import subprocess
from pathlib import Path

HERE = Path(__file__).resolve().parent
SRC = HERE / "csrc" / "oracle_cpu.c"
LIB = HERE / "csrc" / "liboracle_cpu.so"

# x86-64-v3 (AVX2+FMA) rather than -march=native: the .so built in the build container
# travels to the GPU box, whose host CPU may differ.
CFLAGS = ["-O3", "-march=x86-64-v3", "-fopenmp", "-shared", "-fPIC", "-std=c11", "-fno-math-errno"]


def build(force: bool = False) -> Path:
    if LIB.exists() and not force and LIB.stat().st_mtime >= SRC.stat().st_mtime:
        return LIB
    gcc = shutil.which("gcc")
    if gcc is None:
        raise RuntimeError("gcc not found: cannot build the oracle's C restatement")
    cmd = [gcc, *CFLAGS, "-o", str(LIB), str(SRC), "-lm"]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError(f"gcc failed:\n{proc.stdout}\n{proc.stderr}")
    return LIB


def build_ref() -> None:
    """See module docstring: the reference's native code cannot be compiled on Linux."""
    return None


_lib = None


def load() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not LIB.exists():
            build()
        try:
            _lib = ctypes.CDLL(str(LIB))
        except OSError:
            build(force=True)
            _lib = ctypes.CDLL(str(LIB))
    return _lib


if __name__ == "__main__":
    print(build(force=True))
