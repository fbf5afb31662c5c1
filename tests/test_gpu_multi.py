"""Multi-GPU checks as ``-m gpu`` tests: each spawns the torchrun script on N B200s of this
box and self-skips when fewer are visible (the driver's single-GPU tier skips them all)."""

from __future__ import annotations

import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _torchrun(script: str, n: int, port: int, *extra: str, timeout: int = 600) -> str:
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, script), *extra]
    proc = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=ROOT)
    out = proc.stdout + proc.stderr
    assert proc.returncode == 0, out[-4000:]
    return out


@pytest.mark.parametrize("n", [2, 4, 8])
def test_nvlink_allreduce_ranks(n):
    if torch.cuda.device_count() < n:
        pytest.skip(f"needs {n} GPUs, {torch.cuda.device_count()} visible")
    out = _torchrun("tests/multi_gpu_allreduce.py", n, 29540 + n)
    assert "ALLREDUCE OK" in out, out[-4000:]
    print(out[-1500:])
