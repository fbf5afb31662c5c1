"""CPU oracle: a decoder-only forward over the paged path.  TEST INFRASTRUCTURE ONLY.

The attention block follows ``vllm_metal/attention/impls/sdpa.py:463-739``
(``sdpa_forward``: q/k/v projections :297-343, optional per-head q/k RMSNorm
:348-353, RoPE with per-segment offsets :360-368 /
``varlen_rope_compat.py:70-82``, ``reshape_and_cache`` :630-636, the paged
attention primitive :716-732, ``o_proj`` :739).  The layer wiring around it
(pre-norm residual blocks, SwiGLU MLP, final norm, (tied) lm_head, greedy
argmax ``v1/sampling_batch.py:342-348``) is third-party ``mlx_lm`` model code
(mlx-lm @ 254d153, ``mlx_lm/models/llama.py`` / ``qwen3.py``), restated here
from its published structure: **parity unpinned** for that wiring.

All tensors are float32 arrays holding values representable in ``dtype``; every
op rounds once to ``dtype`` where the product stores a tensor (after each
linear, norm, RoPE, activation and residual add), fp32 in between.
"""

from __future__ import annotations

import numpy as np

from . import affine_quant, kv_write, metadata, paged_attention
from ._numeric import round_to


class QWeight:
    """MLX-affine triple + cached fp32 dequantisation (``s*q+b``, unrounded)."""

    def __init__(self, w_q, scales, biases, group_size, bits):
        self.w_q, self.scales, self.biases = w_q, scales, biases
        self.group_size, self.bits = group_size, bits
        self._dense = None

    @property
    def dense(self) -> np.ndarray:
        if self._dense is None:
            self._dense = affine_quant.dequantize(self.w_q, self.scales, self.biases, self.group_size,
                                                  self.bits, "float32")
        return self._dense

    def matmul(self, x: np.ndarray, dtype: str) -> np.ndarray:
        """mx.quantized_matmul(x, w, s, b, transpose=True) -- mlx_native.py:212-222."""
        return round_to(x.astype(np.float32) @ self.dense.T, dtype)

    def rows(self, ids: np.ndarray, dtype: str) -> np.ndarray:
        """mx.dequantize(w[ids], s[ids], b[ids]) -- mlx_native.py:232-239."""
        ids = np.asarray(ids, np.int64)
        return affine_quant.dequantize(self.w_q[ids], self.scales[ids], self.biases[ids], self.group_size,
                                       self.bits, dtype)


def rms_norm(x, weight, eps, dtype):
    """mx.fast.rms_norm: x * rsqrt(mean(x^2) + eps) * w, fp32 inside."""
    xf = x.astype(np.float32)
    inv = 1.0 / np.sqrt((xf * xf).mean(axis=-1, keepdims=True, dtype=np.float32) + np.float32(eps))
    return round_to(xf * inv.astype(np.float32) * weight.astype(np.float32), dtype)


def rope(x, positions, rot_dim, base, traditional, dtype):
    """mlx.nn.RoPE on [T, H, D] at per-token positions (varlen_rope_compat.py:70-82)."""
    t, h, d = x.shape
    half = rot_dim // 2
    inv_freq = np.float64(base) ** (-np.arange(0, half, dtype=np.float64) * 2.0 / rot_dim)
    ang = np.asarray(positions, np.float64)[:, None] * inv_freq[None]
    cs = np.cos(ang)[:, None, :].astype(np.float32)
    sn = np.sin(ang)[:, None, :].astype(np.float32)
    out = x.astype(np.float32).copy()
    if traditional:
        a, b = out[..., 0:rot_dim:2].copy(), out[..., 1:rot_dim:2].copy()
        out[..., 0:rot_dim:2] = a * cs - b * sn
        out[..., 1:rot_dim:2] = b * cs + a * sn
    else:
        a, b = out[..., :half].copy(), out[..., half:rot_dim].copy()
        out[..., :half] = a * cs - b * sn
        out[..., half:rot_dim] = b * cs + a * sn
    return round_to(out, dtype)


def silu_mul(gate_up, dtype):
    inter = gate_up.shape[-1] // 2
    g, u = gate_up[..., :inter].astype(np.float32), gate_up[..., inter:].astype(np.float32)
    s = round_to(g / (1.0 + np.exp(-g)), dtype)
    return round_to(s * u, dtype)


class OracleDecoder:
    """``weights`` is a dict:
    ``embed`` QWeight, ``lm_head`` QWeight|None (tied), ``norm`` [hidden], ``layers`` list of dicts
    with ``input_norm``, ``post_norm`` (or None), ``qkv`` QWeight (rows q|k|v), ``o`` QWeight,
    ``q_norm``/``k_norm`` [D] or None, ``gate_up`` QWeight|None, ``down`` QWeight|None,
    ``sinks`` [H_q] or None, ``window`` int."""

    def __init__(self, weights, n_heads, n_kv_heads, head_dim, rms_eps, rope_base, num_blocks,
                 block_size, dtype="bfloat16", rope_traditional=False, c_attention=False):
        self.w = weights
        self.hq, self.hkv, self.d = n_heads, n_kv_heads, head_dim
        self.eps, self.rope_base, self.trad = rms_eps, rope_base, rope_traditional
        self.dtype = dtype
        self.block_size = block_size
        # c_attention: cache write + attention through the C restatement (oracle/csrc/oracle_cpu.c, checked
        # against the NumPy kernel-exact oracle in tests/test_oracle_golden.py) -- same arithmetic, OpenMP speed.
        self.c_attention = bool(c_attention) and dtype == "bfloat16"
        n_layers = len(weights["layers"])
        if self.c_attention:
            self.kc = [np.zeros((num_blocks, block_size, n_kv_heads, head_dim), np.uint16) for _ in range(n_layers)]
            self.vc = [np.zeros((num_blocks, block_size, n_kv_heads, head_dim), np.uint16) for _ in range(n_layers)]
            return
        self.kc = [np.zeros((num_blocks, block_size, n_kv_heads, head_dim), np.float32) for _ in range(n_layers)]
        self.vc = [np.zeros((num_blocks, block_size, n_kv_heads, head_dim), np.float32) for _ in range(n_layers)]

    def forward(self, input_ids, decode_requests, prefill_requests):
        """One packed step (decode rows first, then prefill; context.py:162-283).
        Returns fp32 logits (rounded to dtype) of every segment's LAST row."""
        dt = self.dtype
        ctx = metadata.prepare_unified(decode_requests, prefill_requests, self.block_size)
        cu = np.asarray(ctx["cu_seqlens"], np.int32)
        seq_lens = np.asarray(ctx["context_lens"], np.int32)
        slots = np.asarray(ctx["slot_mapping"], np.int64)
        bt, kbs = metadata.build_block_tables(ctx["block_tables"], self.block_size)
        assert kbs == self.block_size
        lens = np.diff(cu)
        positions = np.arange(cu[-1]) + np.repeat(np.asarray(ctx["offsets"]) - cu[:-1], lens)
        t = int(cu[-1])
        h = self.w["embed"].rows(np.asarray(input_ids), dt)
        delta = None
        for li, L in enumerate(self.w["layers"]):
            if delta is not None:
                h = round_to(h + delta, dt)
            x = rms_norm(h, L["input_norm"], self.eps, dt)
            qkv = L["qkv"].matmul(x, dt)
            nq, nk = self.hq * self.d, self.hkv * self.d
            q = qkv[:, :nq].reshape(t, self.hq, self.d)
            k = qkv[:, nq:nq + nk].reshape(t, self.hkv, self.d)
            v = qkv[:, nq + nk:].reshape(t, self.hkv, self.d)
            if L.get("q_norm") is not None:
                q = rms_norm(q, L["q_norm"], self.eps, dt)
                k = rms_norm(k, L["k_norm"], self.eps, dt)
            q = rope(q, positions, self.d, self.rope_base, self.trad, dt)
            k = rope(k, positions, self.d, self.rope_base, self.trad, dt)
            if self.c_attention:
                from . import cport
                from ._numeric import bf16_bits, bf16_from_bits

                cport.reshape_and_cache(bf16_bits(k), bf16_bits(v), self.kc[li], self.vc[li], slots)
                a = bf16_from_bits(cport.paged_attention(
                    bf16_bits(q), self.kc[li], self.vc[li], bt, seq_lens, cu, self.d ** -0.5,
                    L.get("window", -1), 0.0, L.get("sinks")))
            else:
                kv_write.reshape_and_cache(k, v, self.kc[li], self.vc[li], slots)
                a = paged_attention.paged_attention_kernel_exact(
                    q, self.kc[li], self.vc[li], bt, seq_lens, cu, self.d ** -0.5, L.get("window", -1), 0.0,
                    L.get("sinks"), dtype=dt)
            a = L["o"].matmul(a.reshape(t, nq), dt)
            if L.get("gate_up") is None and L.get("moe") is None:
                delta = a
                continue
            h = round_to(h + a, dt)
            x = rms_norm(h, L["post_norm"], self.eps, dt)
            if L.get("moe") is not None:
                from . import moe as omoe

                m = L["moe"]
                forced = m.get("forced")  # tests: the product's (indices, scores) of this forward, see gpu_helpers
                ind, sc = forced.pop(0) if forced else (None, None)
                delta, _, _ = omoe.moe_block(x, m["router_w"], m["router_b"], m["gate"], m["up"], m["down"], m["top_k"],
                                             m["group_size"], m["bits"], dt, m["routing"], m["act"], m["norm_topk_prob"],
                                             indices=ind, scores=sc)
                continue
            delta = L["down"].matmul(silu_mul(L["gate_up"].matmul(x, dt), dt), dt)
        if delta is not None:
            h = round_to(h + delta, dt)
        x = rms_norm(h, self.w["norm"], self.eps, dt)
        last = cu[1:] - 1
        head = self.w["lm_head"] if self.w.get("lm_head") is not None else self.w["embed"]
        return head.matmul(x[last], dt)

    @staticmethod
    def greedy(logits):
        """mx.argmax(logits, axis=-1): first maximum (sampling_batch.py:342-348)."""
        return np.argmax(logits, axis=-1).astype(np.int64)
