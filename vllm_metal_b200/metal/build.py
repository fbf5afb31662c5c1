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


def _file_hash(src: Path, extra: list[str]) -> str:
    h = hashlib.sha256()
    h.update(src.read_bytes())
    for f in sorted(list(CSRC_DIR.glob("*.cuh")) + list(INCLUDE_DIR.glob("*.h"))):
        h.update(f.read_bytes())
    h.update(" ".join(NVCC_FLAGS + extra).encode())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every ``csrc/*.cu`` to an object (in parallel, cached by content hash under
    ``build/``) and link them into the one shared library."""
    from concurrent.futures import ThreadPoolExecutor

    out = output_path()
    if not force and out.exists() and not stale_artifacts():
        return out
    extra = os.environ.get("B200_NVCC_EXTRA", "").split()  # tuning experiments (-DQGEMM_...), not for shipping
    nvcc = find_nvcc()
    obj_dir = _PKG_DIR / "build"
    obj_dir.mkdir(exist_ok=True)
    compile_flags = [f for f in NVCC_FLAGS if f != "--shared"]
    logs: list[str] = []

    def compile_one(src: Path) -> Path:
        obj = obj_dir / (src.stem + ".o")
        tag = obj_dir / (src.stem + ".hash")
        hv = _file_hash(src, extra)
        if not force and obj.exists() and tag.exists() and tag.read_text() == hv:
            return obj
        cmd = [nvcc, *compile_flags, *extra, "-c", "-o", str(obj), str(src)]
        proc = subprocess.run(cmd, capture_output=True, text=True)
        logs.append(" ".join(cmd) + "\n" + proc.stdout + proc.stderr)
        if proc.returncode != 0:
            raise RuntimeError(f"nvcc failed ({proc.returncode}) on {src.name}:\n{(proc.stdout + proc.stderr)[-4000:]}")
        tag.write_text(hv)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        objs = list(ex.map(compile_one, sources()))
    cmd = [nvcc, "--shared", "-o", str(out), *[str(o) for o in objs]]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    logs.append(" ".join(cmd) + "\n" + proc.stdout + proc.stderr)
    (_PKG_DIR / "build.log").write_text("\n".join(logs))
    if proc.returncode != 0:
        raise RuntimeError(f"link failed ({proc.returncode}):\n{(proc.stdout + proc.stderr)[-4000:]}")
    if verbose:
        print("\n".join(logs))
    stamp_path().write_text(source_hash() + "\n")
    return out


if __name__ == "__main__":
    p = build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(p)
