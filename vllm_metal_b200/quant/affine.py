"""MLX affine group quantisation in torch (load-time helper).

Restates ``mx.quantize`` (MLX 0.32, third-party) so
random-init or dense checkpoints can be turned into the MLX triple
``weight`` (uint32 bit patterns stored as int32) / ``scales`` / ``biases``.
Runs on whatever device the weight lives on; it is load-time code, not part of
the decode step.  Bit-exact against ``oracle.affine_quant.quantize`` on CPU
(tests/test_quant_host.py).
"""

from __future__ import annotations

import torch


def pack(q: torch.Tensor, bits: int) -> torch.Tensor:
    """uint8-valued ``[N, K]`` -> int32 ``[N, K*bits/32]`` (element k at bit (k % per)*bits)."""
    per = 32 // bits
    n, k = q.shape
    u = q.to(torch.int64).reshape(n, k // per, per)
    shifts = torch.arange(per, device=q.device, dtype=torch.int64) * bits
    word = (u << shifts).sum(dim=-1)  # disjoint bit fields: sum == or
    word = torch.where(word >= 2**31, word - 2**32, word)
    return word.to(torch.int32)


def unpack(w: torch.Tensor, bits: int) -> torch.Tensor:
    """int32 ``[N, K*bits/32]`` -> uint8 ``[N, K]``."""
    per = 32 // bits
    u = w.to(torch.int64) & 0xFFFFFFFF
    shifts = torch.arange(per, device=w.device, dtype=torch.int64) * bits
    return ((u[:, :, None] >> shifts) & ((1 << bits) - 1)).reshape(w.shape[0], -1).to(torch.uint8)


def quantize(w: torch.Tensor, group_size: int = 64, bits: int = 4, dtype: torch.dtype = torch.bfloat16):
    """Returns ``(weight int32 [N, K*bits/32], scales [N, K/g], biases [N, K/g])``."""
    if group_size not in (32, 64, 128) or bits not in (4, 8):
        raise ValueError(f"unsupported quantisation group_size={group_size} bits={bits}")
    n, k = w.shape
    if k % group_size:
        raise ValueError(f"in_features {k} not divisible by group_size {group_size}")
    n_bins = float((1 << bits) - 1)
    wg = w.float().reshape(n, k // group_size, group_size)
    w_max = wg.amax(dim=-1)
    w_min = wg.amin(dim=-1)
    side = w_min.abs() > w_max.abs()
    s0 = torch.clamp((w_max - w_min) / n_bins, min=1e-7)
    s = torch.where(side, s0, -s0)
    edge = torch.where(side, w_min, w_max)
    q0 = torch.round(edge / s)
    nz = q0 != 0
    s = torch.where(nz, edge / torch.where(nz, q0, torch.ones_like(q0)), s)
    b = torch.where(nz, edge, torch.zeros_like(edge))
    s = s.to(dtype)
    b = b.to(dtype)
    q = torch.clamp(torch.round((wg - b.float()[..., None]) / s.float()[..., None]), 0, n_bins)
    return pack(q.reshape(n, k).to(torch.uint8), bits), s, b


def dequantize_reference(w, scales, biases, group_size: int = 64, bits: int = 4) -> torch.Tensor:
    """fp32 ``s*q+b`` in plain torch -- load-time / debugging only."""
    q = unpack(w, bits).float()
    s = scales.float().repeat_interleave(group_size, dim=1)
    b = biases.float().repeat_interleave(group_size, dim=1)
    return s * q + b
