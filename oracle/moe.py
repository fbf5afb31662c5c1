"""CPU oracle: quantized mixture-of-experts MLP.  TEST INFRASTRUCTURE ONLY.

Restates mlx-lm's MoE block as the reference reaches it for gpt-oss / Qwen3-MoE checkpoints
(``QuantizedSwitchLinear`` named in vllm_metal/quant/awq_loader.py:171-173): router logits, top-k selection, softmax
over the selected logits (gpt-oss) or over all experts followed by top-k and renormalisation (Qwen3-MoE), per-expert
``mx.gather_qmm`` projections (= ``affine_quant.quantized_matmul`` with that expert's weights), SwiGLU (plain, or
gpt-oss's clipped form: ``g = min(gate, 7); u = clip(up, +-7); g * sigmoid(1.702 g) * (u + 1)``), score-weighted sum.

The arithmetic lives in third-party code that is NOT under /root/reference (mlx 0.32.0 ``gather_qmm``, mlx-lm 254d153
``switch_layers.py`` / ``gpt_oss.py``), and no reference test holds a vector for it: **parity unpinned**.
Precision policy as everywhere in the oracle: fp32 arithmetic on dtype-rounded inputs, one rounding per stored tensor.
"""

from __future__ import annotations

import numpy as np

from . import affine_quant as oq
from ._numeric import round_to


def softmax(x: np.ndarray) -> np.ndarray:
    x = x.astype(np.float32)
    e = np.exp(x - x.max(axis=-1, keepdims=True))
    return (e / e.sum(axis=-1, keepdims=True)).astype(np.float32)


def route(logits: np.ndarray, top_k: int, routing: str = "topk_softmax", norm_topk_prob: bool = True):
    """-> (indices [T, k] by descending score, lowest expert id first on ties; scores f32 [T, k])."""
    logits = logits.astype(np.float32)
    if routing == "topk_softmax":
        idx = np.argsort(-logits, axis=-1, kind="stable")[:, :top_k]
        return idx, softmax(np.take_along_axis(logits, idx, axis=-1))
    probs = softmax(logits)
    idx = np.argsort(-probs, axis=-1, kind="stable")[:, :top_k]
    vals = np.take_along_axis(probs, idx, axis=-1)
    if norm_topk_prob:
        vals = vals / vals.sum(axis=-1, keepdims=True)
    return idx, vals.astype(np.float32)


def activation(gate: np.ndarray, up: np.ndarray, kind: str = "silu") -> np.ndarray:
    g, u = gate.astype(np.float32), up.astype(np.float32)
    if kind == "silu":
        return g / (1.0 + np.exp(-g)) * u
    if kind == "gpt_oss":
        g = np.minimum(g, np.float32(7.0))
        u = np.clip(u, np.float32(-7.0), np.float32(7.0))
        return g / (1.0 + np.exp(-np.float32(1.702) * g)) * (u + np.float32(1.0))
    raise ValueError(kind)


def switch_linear(x_rows: np.ndarray, experts: np.ndarray, w, s, b, bias, group_size: int, bits: int, dtype: str):
    """``gather_qmm``: row i of ``x_rows`` times expert ``experts[i]``'s weight.  ``w`` uint32 [E, N, K*bits/32]."""
    out = np.zeros((x_rows.shape[0], w.shape[1]), np.float32)
    for e in np.unique(experts):
        rows = np.nonzero(experts == e)[0]
        out[rows] = oq.quantized_matmul(x_rows[rows], w[e], s[e], b[e], group_size, bits,
                                        None if bias is None else bias[e], dtype)
    return out


def moe_block(x, router_w, router_b, gate, up, down, top_k, group_size=64, bits=4, dtype="bfloat16",
              routing="topk_softmax", act="silu", norm_topk_prob=True, indices=None, scores=None):
    """``x`` f32 [T, hidden] (dtype-representable); ``router_w`` dense f32 [E, hidden] (+ ``router_b``);
    ``gate`` / ``up`` / ``down``: tuples ``(w, scales, biases, bias|None)`` of stacked expert tensors.
    ``indices`` / ``scores`` override the routing (to compare the expert path on the product's own routing)."""
    t = x.shape[0]
    if indices is None:
        logits = round_to(x.astype(np.float32) @ router_w.astype(np.float32).T +
                          (0 if router_b is None else router_b.astype(np.float32)), dtype)
        indices, scores = route(logits, top_k, routing, norm_topk_prob)
    k = indices.shape[1]
    flat = indices.reshape(-1)
    xr = np.repeat(x, k, axis=0)
    g = switch_linear(xr, flat, *gate, group_size, bits, dtype)
    u = switch_linear(xr, flat, *up, group_size, bits, dtype)
    h = round_to(activation(g, u, act), dtype)
    y = switch_linear(h, flat, *down, group_size, bits, dtype).reshape(t, k, -1)
    return round_to((y * scores[..., None].astype(np.float32)).sum(axis=1), dtype), indices, scores
