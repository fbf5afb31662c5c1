"""Build the sm_100a shared library behind ``get_ops()``.

Reference counterpart: ``vllm_metal/metal/build.py`` (clang++ + ``xcrun metal``
with staleness stamps, :79-91,236-306).  Here one ``nvcc`` invocation compiles
every ``csrc/*.cu`` for ``sm_100a`` into ``lib_b200_paged_ops.so`` next to the
package; a stamp file records the source hash so ``stale_artifacts()`` can tell
a developer that sources changed without a rebuild.  ``python -m
vllm_metal_b200.metal.build`` rebuilds.
"""

from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

_PKG_DIR = Path(__file__).resolve().parent.parent
CSRC_DIR = _PKG_DIR / "csrc"
INCLUDE_DIR = _PKG_DIR.parent / "include"
LIB_NAME = "lib_b200_paged_ops.so"

NVCC_FLAGS = [
    "-gencode",
    "arch=compute_100a,code=sm_100a",
    "-lineinfo",
    "-O3",
    "-std=c++17",
    "--shared",
    "-Xcompiler",
    "-fPIC",
    "-Xptxas",
    "-v",
    # IEEE div/exp like the reference's -fno-fast-math shader build (build.py:69-76)
]


def output_path() -> Path:
    return _PKG_DIR / LIB_NAME


def stamp_path() -> Path:
    return _PKG_DIR / (LIB_NAME + ".stamp")


def sources() -> list[Path]:
    return sorted(CSRC_DIR.glob("*.cu"))


def source_hash() -> str:
    h = hashlib.sha256()
    files = sorted(list(CSRC_DIR.glob("*.cu")) + list(CSRC_DIR.glob("*.cuh")) + list(INCLUDE_DIR.glob("*.h")))
    for f in files:
        h.update(f.name.encode())
        h.update(f.read_bytes())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def stale_artifacts() -> list[Path]:
    """Artifacts whose stamp no longer matches the sources (empty if no sources
    are present, e.g. on a box that only received the built library)."""
    if not output_path().exists():
        return [output_path()]
    if not stamp_path().exists() or not CSRC_DIR.exists():
        return []
    return [] if stamp_path().read_text().strip() == source_hash() else [output_path()]


def find_nvcc() -> str:
    cand = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not Path(cand).exists():
        raise RuntimeError("nvcc not found; set NVCC or install the CUDA toolkit")
    return cand


def build(force: bool = False, verbose: bool = False) -> Path:
    out = output_path()
    if not force and out.exists() and not stale_artifacts():
        return out
    extra = os.environ.get("B200_NVCC_EXTRA", "").split()  # tuning experiments (-DQGEMM_...), not for shipping
    cmd = [find_nvcc(), *NVCC_FLAGS, *extra, "-o", str(out), *[str(s) for s in sources()]]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    log = proc.stdout + proc.stderr
    (_PKG_DIR / "build.log").write_text(" ".join(cmd) + "\n" + log)
    if proc.returncode != 0:
        raise RuntimeError(f"nvcc failed ({proc.returncode}):\n{log[-4000:]}")
    if verbose:
        print(log)
    stamp_path().write_text(source_hash() + "\n")
    return out


if __name__ == "__main__":
    p = build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(p)
