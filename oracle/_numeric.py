"""Small numeric helpers for the oracle (test infrastructure only).

NumPy has no bfloat16, so 16-bit tensors are carried as float32 arrays whose
values are exactly representable in the target dtype; ``round_to`` applies the
rounding a store to that dtype would perform (round-to-nearest-even).
"""

from __future__ import annotations

import numpy as np

DTYPES = ("float32", "float16", "bfloat16")


def bf16_round(x: np.ndarray) -> np.ndarray:
    """Round float32 values to bfloat16 (RNE) and return them as float32."""
    a = np.ascontiguousarray(x, dtype=np.float32)
    u = a.view(np.uint32).astype(np.uint64)
    lsb = (u >> 16) & 1
    r = ((u + 0x7FFF + lsb) >> 16) << 16
    out = r.astype(np.uint32).view(np.float32).reshape(a.shape)
    nan = np.isnan(a)
    if nan.any():
        out = out.copy()
        out[nan] = np.nan
    return out


def round_to(x: np.ndarray, dtype: str) -> np.ndarray:
    """Round ``x`` to ``dtype`` and return float32 holding the rounded values."""
    if dtype == "float32":
        return np.asarray(x, dtype=np.float32)
    if dtype == "float16":
        return np.asarray(x, dtype=np.float32).astype(np.float16).astype(np.float32)
    if dtype == "bfloat16":
        return bf16_round(x)
    raise ValueError(f"unsupported dtype {dtype!r}; expected one of {DTYPES}")


def bf16_bits(x: np.ndarray) -> np.ndarray:
    """uint16 bit patterns of bf16-representable float32 values."""
    a = np.ascontiguousarray(x, dtype=np.float32)
    return (a.view(np.uint32) >> 16).astype(np.uint16).reshape(a.shape)


def bf16_from_bits(b: np.ndarray) -> np.ndarray:
    """float32 values from bf16 bit patterns (uint16)."""
    u = np.ascontiguousarray(b, dtype=np.uint16).astype(np.uint32) << 16
    return u.view(np.float32).reshape(b.shape)
