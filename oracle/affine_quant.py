"""CPU oracle: MLX affine int4/int8 group quantisation.  TEST INFRASTRUCTURE ONLY.

The arithmetic lives in a third-party dependency that is NOT under
``/root/reference``: **mlx == 0.32.0** (``pyproject.toml:36``), functions
``mx.quantize`` / ``mx.dequantize`` / ``mx.quantized_matmul`` and
``mlx.nn.QuantizedLinear``; the AWQ repack lives in **mlx-lm @ 254d153**
(``pyproject.toml:40``, ``mlx_lm.utils._transform_awq_weights``).  MLX is not
installed here, so this file restates MLX's published affine scheme and anchors
on the reference's own call sites and tests:

* call sites: ``vllm_metal/gguf/mlx_native.py:212-222`` (quantized_matmul,
  transpose=True), ``:232-239`` (dequantize rows), ``attention/impls/sdpa.py:297-343,739``.
* shapes: ``weight`` uint32 ``[N, K*bits/32]``, ``scales``/``biases`` ``[N, K/g]``
  (``tests/quant/test_awq_repack_parity.py:143-152``,
  ``vllm_metal/gguf/mlx_native.py:126-133`` "a group of 32 weights is ``bits`` words").
* semantics: ``W' = scale * q + bias`` with ``bias = -(z * s)`` for AWQ imports
  (``tests/quant/test_awq_repack_parity.py:56-60,160-164``).

**parity unpinned** (no reference test fixes them): the bit order inside a
uint32 word (restated: element ``k`` at bits ``[(k % (32/b))*b, +b)`` of word
``k // (32/b)``, lowest bits first), the quantiser's rounding rule and the
accumulation order of ``quantized_matmul``.
"""

from __future__ import annotations

import numpy as np

from ._numeric import round_to

AWQ_PACK_ORDER = (0, 2, 4, 6, 1, 3, 5, 7)  # tests/quant/test_awq_repack_parity.py:31


def pack(q: np.ndarray, bits: int) -> np.ndarray:
    """``q`` uint [N, K] (values < 2**bits) -> uint32 [N, K*bits/32], low bits first."""
    assert bits in (4, 8)
    per = 32 // bits
    n, k = q.shape
    assert k % per == 0
    u = q.astype(np.uint32).reshape(n, k // per, per)
    shifts = (np.arange(per, dtype=np.uint32) * np.uint32(bits))[None, None, :]
    return np.bitwise_or.reduce(u << shifts, axis=-1).astype(np.uint32)


def unpack(w: np.ndarray, bits: int) -> np.ndarray:
    """uint32 [N, K*bits/32] -> uint8 [N, K]."""
    assert bits in (4, 8)
    per = 32 // bits
    n, kw = w.shape
    shifts = (np.arange(per, dtype=np.uint32) * np.uint32(bits))[None, None, :]
    mask = np.uint32((1 << bits) - 1)
    return ((w.astype(np.uint32)[:, :, None] >> shifts) & mask).reshape(n, kw * per).astype(np.uint8)


def quantize(w: np.ndarray, group_size: int = 64, bits: int = 4, dtype: str = "bfloat16"):
    """MLX ``mx.quantize`` affine mode, restated.

    Returns ``(w_q uint32 [N, K*bits/32], scales f32 [N, K/g], biases f32 [N, K/g])``
    with scales/biases already rounded to ``dtype``.
    """
    assert group_size in (32, 64, 128) and bits in (4, 8)
    n, k = w.shape
    assert k % group_size == 0
    n_bins = np.float32((1 << bits) - 1)
    wg = w.astype(np.float32).reshape(n, k // group_size, group_size)
    w_max = wg.max(axis=-1)
    w_min = wg.min(axis=-1)
    side = np.abs(w_min) > np.abs(w_max)
    s0 = np.maximum((w_max - w_min) / n_bins, np.float32(1e-7)).astype(np.float32)
    s = np.where(side, s0, -s0).astype(np.float32)
    edge = np.where(side, w_min, w_max).astype(np.float32)
    q0 = np.rint(edge / s).astype(np.float32)
    nz = q0 != 0
    s = np.where(nz, edge / np.where(nz, q0, 1), s).astype(np.float32)
    b = np.where(nz, edge, 0).astype(np.float32)
    s = round_to(s, dtype)
    b = round_to(b, dtype)
    q = np.clip(np.rint((wg - b[..., None]) / s[..., None]), 0, n_bins).astype(np.uint8)
    return pack(q.reshape(n, k), bits), s, b


def dequantize(
    w_q: np.ndarray,
    scales: np.ndarray,
    biases: np.ndarray,
    group_size: int = 64,
    bits: int = 4,
    dtype: str = "float32",
) -> np.ndarray:
    """``mx.dequantize``: ``s * q + b`` in fp32, rounded once to ``dtype``."""
    q = unpack(w_q, bits).astype(np.float32)
    n, k = q.shape
    s = np.repeat(scales.astype(np.float32), group_size, axis=1)
    b = np.repeat(biases.astype(np.float32), group_size, axis=1)
    assert s.shape == (n, k), (s.shape, (n, k))
    return round_to(s * q + b, dtype)


def quantized_matmul(
    x: np.ndarray,
    w_q: np.ndarray,
    scales: np.ndarray,
    biases: np.ndarray,
    group_size: int = 64,
    bits: int = 4,
    bias: np.ndarray | None = None,
    dtype: str = "float32",
) -> np.ndarray:
    """``mx.quantized_matmul(x, w, s, b, transpose=True)`` = ``x @ W'^T`` (+bias).

    fp32 accumulate, result rounded to the activation dtype (mlx_native.py:209-222);
    optional dense bias added after (gguf/wrappers.py:71-75).
    """
    w = dequantize(w_q, scales, biases, group_size, bits, "float32")
    y = x.astype(np.float32) @ w.T
    if bias is not None:
        y = y + bias.astype(np.float32)
    return round_to(y, dtype)


def quantized_matmul_factored(x, w_q, scales, biases, group_size=64, bits=4):
    """The factorisation MLX's kernels use: ``sum_g s_g*sum(x*q) + b_g*sum(x)`` (fp32)."""
    q = unpack(w_q, bits).astype(np.float32)
    n, k = q.shape
    ng = k // group_size
    xg = x.astype(np.float32).reshape(-1, ng, group_size)
    qg = q.reshape(n, ng, group_size)
    xq = np.einsum("mgk,ngk->mng", xg, qg)
    xs = xg.sum(axis=-1)
    return (xq * scales.astype(np.float32)[None] + xs[:, None, :] * biases.astype(np.float32)[None]).sum(-1)


# -- AWQ -> MLX affine repack (mlx_lm._transform_awq_weights, third-party) ----


def awq_unpack_int32(packed: np.ndarray) -> np.ndarray:
    """[..., C] int32 in AutoAWQ interleaved order -> [..., C*8] uint8 nibbles."""
    u = packed.astype(np.uint32)
    out = np.zeros(u.shape[:-1] + (u.shape[-1] * 8,), dtype=np.uint8)
    for shift_idx, pos in enumerate(AWQ_PACK_ORDER):
        out[..., pos::8] = ((u >> np.uint32(shift_idx * 4)) & np.uint32(0xF)).astype(np.uint8)
    return out


def awq_pack_int32(unpacked: np.ndarray) -> np.ndarray:
    """Inverse of ``awq_unpack_int32`` (tests/quant/test_awq_repack_parity.py:34-45)."""
    u = unpacked.astype(np.uint32) & np.uint32(0xF)
    packed = np.zeros(u.shape[:-1] + (u.shape[-1] // 8,), dtype=np.uint32)
    for shift_idx, pos in enumerate(AWQ_PACK_ORDER):
        packed |= u[..., pos::8] << np.uint32(shift_idx * 4)
    return packed.astype(np.int32)


def awq_to_mlx_affine(qweight, qzeros, scales, dtype: str = "float16"):
    """AWQ GEMM triple -> MLX affine triple.

    ``qweight`` [in, out/8] int32, ``qzeros`` [groups, out/8] int32, ``scales``
    [groups, out] -> ``weight`` uint32 [out, in/8], ``scales`` [out, groups],
    ``biases = -(zeros * scales)`` [out, groups]
    (tests/quant/test_awq_repack_parity.py:143-164).
    """
    q = awq_unpack_int32(qweight)  # [in, out]
    z = awq_unpack_int32(qzeros).astype(np.float32)  # [groups, out]
    s = scales.astype(np.float32)
    weight = pack(np.ascontiguousarray(q.T), 4)
    s_t = np.ascontiguousarray(s.T)
    b_t = round_to(-(np.ascontiguousarray(z.T) * s_t), dtype)
    return weight, round_to(s_t, dtype), b_t
