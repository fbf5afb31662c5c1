"""CPU oracle: paged varlen attention.  TEST INFRASTRUCTURE ONLY.

Two restatements of the same path:

``ref_paged_attn``
    Follows the reference's own test oracle line by line
    (``/root/reference/tools/attention_bench_utils.py:11-62``): gather K/V by
    block table, ``q*scale``, fp32 scores, causal ``triu(k=kv-q+1)``, window
    ``not triu(k=kv-(q+W)+1)``, optional softcap, softmax, cast to V dtype, PV.
    Sinks follow the semantics of ``mx.fast.scaled_dot_product_attention(...,
    sinks=)`` as the reference tests use it
    (``tests/test_paged_attention_sinks.py:102-115``): softmax over
    ``[scores, sink]`` with the sink column dropped.

``paged_attention_kernel_exact``
    Follows the Metal kernels where they differ from the test oracle:
    effective context ``ctx - q_len + q_pos + 1`` and "empty rows leave out
    untouched" (``pagedattention.metal:913-926``), mask + log2 domain
    (``:1129-1138``), sink fold (``:1318-1328`` / ``:2045-2072``) and the
    ``1/(l + 1e-6)`` normalisation (``:1383``, ``:2074``,
    ``pagedattention_tiled.metal:496``).  fp32 throughout, one final rounding.

Arrays are float32 holding values representable in ``dtype`` (see
``_numeric``).  Cache layout ``[num_blocks, block_size, H_kv, D]``
(``vllm_metal/attention/caches/kv_cache.py:223-234``).
"""

from __future__ import annotations

import numpy as np

from ._numeric import round_to

LOG2E = np.float32(1.4426950408889634)


def gather_kv(cache: np.ndarray, block_table_row: np.ndarray, kv_len: int) -> np.ndarray:
    """``cache[block_indices].reshape(-1, H, D)[:kv_len]`` (attention_bench_utils.py:30-34)."""
    _, block_size, h, d = cache.shape
    nblk = (kv_len + block_size - 1) // block_size
    idx = np.asarray(block_table_row[:nblk], dtype=np.int64)
    return cache[idx].reshape(-1, h, d)[:kv_len]


def ref_paged_attn(
    query: np.ndarray,
    key_cache: np.ndarray,
    value_cache: np.ndarray,
    query_lens: list[int],
    kv_lens: list[int],
    block_tables: np.ndarray,
    scale: float,
    sliding_window: int | None = None,
    soft_cap: float | None = None,
    sinks: np.ndarray | None = None,
    dtype: str = "float32",
) -> np.ndarray:
    """Restatement of tools/attention_bench_utils.py:11-62 (+ SDPA sink column)."""
    _, _, num_kv_heads, _ = key_cache.shape
    outputs = []
    start = 0
    for i, q_len in enumerate(query_lens):
        kv_len = kv_lens[i]
        # q * scale happens in the query dtype (line 28)
        q = round_to(query[start : start + q_len].astype(np.float32) * np.float32(scale), dtype)
        k = gather_kv(key_cache, block_tables[i], kv_len)
        v = gather_kv(value_cache, block_tables[i], kv_len)
        if q.shape[1] != k.shape[1]:
            n_rep = q.shape[1] // k.shape[1]
            k = np.repeat(k, n_rep, axis=1)
            v = np.repeat(v, n_rep, axis=1)
        # einsum in the tensor dtype, then astype(float32) (line 41)
        attn = round_to(np.einsum("qhd,khd->hqk", q, k, dtype=np.float32), dtype)
        empty = np.ones((q_len, kv_len), dtype=np.float32)
        mask = np.triu(empty, k=kv_len - q_len + 1).astype(bool)
        if sliding_window is not None:
            sw = np.logical_not(
                np.triu(empty, k=kv_len - (q_len + sliding_window) + 1).astype(bool)
            )
            mask = np.logical_or(mask, sw)
        if soft_cap is not None and soft_cap > 0:
            attn = np.float32(soft_cap) * np.tanh(attn / np.float32(soft_cap))
        attn = np.where(mask[None], -np.inf, attn).astype(np.float32)
        if sinks is not None:
            col = np.broadcast_to(
                np.asarray(sinks, np.float32)[:, None, None], (attn.shape[0], q_len, 1)
            )
            attn = np.concatenate([attn, col], axis=-1)
        m = attn.max(axis=-1, keepdims=True)
        m = np.where(np.isfinite(m), m, 0.0)
        e = np.exp(attn - m)
        p = e / e.sum(axis=-1, keepdims=True)
        if sinks is not None:
            p = p[..., :-1]
        p = round_to(p, dtype)  # .astype(v.dtype) (line 58)
        out = np.einsum("hqk,khd->qhd", p, v, dtype=np.float32)
        outputs.append(round_to(out, dtype))
        start += q_len
    return np.concatenate(outputs, axis=0)


def paged_attention_kernel_exact(
    query: np.ndarray,
    key_cache: np.ndarray,
    value_cache: np.ndarray,
    block_tables: np.ndarray,
    seq_lens: np.ndarray,
    cu_seqlens_q: np.ndarray,
    scale: float,
    sliding_window: int = -1,
    softcap: float = 0.0,
    sinks: np.ndarray | None = None,
    out: np.ndarray | None = None,
    dtype: str = "float32",
) -> np.ndarray:
    """Restatement of the Metal kernels' arithmetic in fp32 (see module docstring).

    ``out`` (if given) is updated in place and rows with no KV to attend to are
    left untouched, as pagedattention.metal:917-920 does.
    """
    t_total, h_q, d = query.shape
    _, block_size, h_kv, _ = key_cache.shape
    g = h_q // h_kv
    if out is None:
        out = np.zeros((t_total, h_q, d), dtype=np.float32)
    num_seqs = len(cu_seqlens_q) - 1
    for s in range(num_seqs):
        q0, q1 = int(cu_seqlens_q[s]), int(cu_seqlens_q[s + 1])
        q_len = q1 - q0
        ctx = int(seq_lens[s])
        if q_len <= 0:
            continue
        k_all = gather_kv(key_cache, block_tables[s], ctx).astype(np.float32)
        v_all = gather_kv(value_cache, block_tables[s], ctx).astype(np.float32)
        for qp in range(q_len):
            eff = ctx - q_len + qp + 1  # pagedattention.metal:913
            if eff <= 0:
                continue
            lo = 0
            if sliding_window >= 0:
                lo = max(0, eff - sliding_window)  # :1131-1133
            for h in range(h_q):
                kvh = h // g  # :959-960
                qv = query[q0 + qp, h].astype(np.float32)
                sc = (k_all[lo:eff, kvh] @ qv) * np.float32(scale)
                if softcap > 0.0:
                    sc = np.tanh(sc / np.float32(softcap)) * np.float32(softcap)
                sc = sc.astype(np.float32) * LOG2E  # :1138
                m = sc.max() if sc.size else np.float32(-np.inf)
                if sinks is not None:
                    m = max(m, np.float32(sinks[h]) * LOG2E)  # :1318-1328
                if not np.isfinite(m):
                    m = np.float32(0.0)
                w = np.exp2(sc - m).astype(np.float32)
                l = w.sum(dtype=np.float32)
                if sinks is not None:
                    l = l + np.exp2(np.float32(sinks[h]) * LOG2E - m)
                o = w @ v_all[lo:eff, kvh]
                out[q0 + qp, h] = o / (l + np.float32(1e-6))  # :1383
    return round_to(out, dtype)
