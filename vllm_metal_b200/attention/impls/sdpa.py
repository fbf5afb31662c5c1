"""SDPA forward over the paged cache: project -> norm -> RoPE -> cache write ->
paged attention -> o_proj.

Host-side mirror of the reference's ``vllm_metal/attention/impls/sdpa.py``
(``_pick_kernel_block_size`` :115-124, ``_build_block_tables`` :127-163,
``_kernel_metadata`` :181-211, ``sdpa_forward`` :463-739) with torch tensors
and the C-ABI ops.  The attention module contract is the mlx_lm one the
reference duck-types (``q_proj/k_proj/v_proj/o_proj``, ``n_heads``,
``n_kv_heads``, ``scale`` or ``sm_scale``, optional ``q_norm``/``k_norm``,
``rope``, optional ``sinks``).
"""

from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import torch

from ...metal import get_ops
from ...metal.constants import KERNEL_BLOCK_SIZES
from ..caches.kv_cache import B200PagedKVCache
from ..context import PagedAttentionContext


# Test knob: False runs the reference's step-by-step sequence (norm, rope, reshape_and_cache as
# separate launches) instead of the fused kernel; both are bit-identical (tests/test_gpu_model.py).
_FUSE_NORM_ROPE_CACHE = True


def _pick_kernel_block_size(cache_block_size: int) -> int:
    """Largest kernel page size dividing the cache block size (sdpa.py:115-124)."""
    for kbs in KERNEL_BLOCK_SIZES:
        if cache_block_size % kbs == 0:
            return kbs
    raise ValueError(
        f"Cache block_size={cache_block_size} is not divisible by any "
        f"supported kernel block size {KERNEL_BLOCK_SIZES}. "
        "Adjust --block-size (must be a multiple of 8)."
    )


def _build_block_tables_np(raw_block_tables: list[list[int]], cache_block_size: int) -> tuple[np.ndarray, int]:
    """Host int32 table ``[S, max_blocks(*ratio)]`` + kernel page size.

    Rows are padded with 0; a cache block size outside {32,16,8} expands block
    ``b`` into ``b*r .. b*r+r-1`` with ``r = cache_block_size // kernel_block_size``
    (sdpa.py:127-163); the cache is viewed with the kernel page size to match.
    """
    if not raw_block_tables:
        return np.zeros((0, 0), dtype=np.int32), cache_block_size
    width = max(len(row) for row in raw_block_tables)
    table = np.zeros((len(raw_block_tables), width), dtype=np.int32)
    for i, row in enumerate(raw_block_tables):
        table[i, : len(row)] = row
    if cache_block_size in KERNEL_BLOCK_SIZES:
        return table, cache_block_size
    kbs = _pick_kernel_block_size(cache_block_size)
    ratio = cache_block_size // kbs
    expanded = table[:, :, None] * ratio + np.arange(ratio, dtype=np.int32)
    return expanded.reshape(table.shape[0], -1).astype(np.int32), kbs


def _build_block_tables(raw_block_tables, cache_block_size, device="cuda"):
    """Device version of :func:`_build_block_tables_np` (reference return shape)."""
    table, kbs = _build_block_tables_np(raw_block_tables, cache_block_size)
    return _to_device(table, torch.int32, device), kbs


def _to_device(arr: np.ndarray, dtype: torch.dtype, device) -> torch.Tensor:
    """Pinned staging + async H2D on the current stream."""
    host = torch.from_numpy(np.ascontiguousarray(arr)).to(dtype)
    if host.numel() and torch.cuda.is_available():
        host = host.pin_memory()
    return host.to(device, non_blocking=True)


@dataclass(frozen=True, eq=False)
class _KernelMetadata:
    """Device copies of the per-forward paged metadata for one KV group."""

    slot_mapping: torch.Tensor  # int64 [T]
    seq_lens: torch.Tensor  # int32 [S]
    cu_seqlens_q: torch.Tensor  # int32 [S+1]
    block_tables: torch.Tensor  # int32 [S, max_blocks]
    block_size: int
    max_seq_len: int
    positions: torch.Tensor  # int32 [T] RoPE position of every packed token


def _positions(ctx: PagedAttentionContext) -> np.ndarray:
    cu = np.asarray(ctx.cu_seqlens, dtype=np.int64)
    lens = np.diff(cu)
    starts = np.repeat(np.asarray(ctx.offsets, dtype=np.int64) - cu[:-1], lens)
    return (np.arange(cu[-1], dtype=np.int64) + starts).astype(np.int32)


def _kernel_metadata(ctx, group_index, raw_slot_mapping, raw_block_tables, cache_block_size, device="cuda"):
    """Built on the group's first layer, reused by the rest (sdpa.py:181-211)."""
    key = (group_index, cache_block_size)
    meta = ctx.kernel_metadata_cache.get(key)
    if meta is None:
        table, kbs = _build_block_tables_np(raw_block_tables, cache_block_size)
        meta = _KernelMetadata(
            slot_mapping=_to_device(np.asarray(raw_slot_mapping, dtype=np.int64), torch.int64, device),
            seq_lens=_to_device(np.asarray(ctx.context_lens, dtype=np.int32), torch.int32, device),
            cu_seqlens_q=_to_device(np.asarray(ctx.cu_seqlens, dtype=np.int32), torch.int32, device),
            block_tables=_to_device(table, torch.int32, device),
            block_size=kbs,
            max_seq_len=max(ctx.context_lens) if ctx.context_lens else 0,
            positions=_to_device(_positions(ctx), torch.int32, device),
        )
        ctx.kernel_metadata_cache[key] = meta
    return meta


def sdpa_forward(
    inner,
    x: torch.Tensor,
    ctx: PagedAttentionContext,
    kv_cache: B200PagedKVCache,
    layer_idx: int,
    *,
    meta: _KernelMetadata | None = None,
) -> torch.Tensor:
    """One attention layer over the packed batch ``x [T, hidden]`` (or ``[1, T, hidden]``).

    Steps follow sdpa.py:463-739: quantized q/k/v projections, optional per-head
    q/k RMSNorm, RoPE at per-token positions, ``reshape_and_cache``,
    ``paged_attention_primitive`` (softcap 0.0, per-layer sliding window,
    optional sinks), ``o_proj``.  ``meta`` lets a graph-captured runner pass
    persistent device metadata instead of the per-step context.
    """
    ops = get_ops()
    lead_shape = x.shape[:-1]
    x2 = x.reshape(-1, x.shape[-1])
    t = x2.shape[0]
    n_heads = getattr(inner, "n_heads", None) or inner.num_attention_heads
    n_kv_heads = getattr(inner, "n_kv_heads", None) or inner.num_key_value_heads
    scale = getattr(inner, "scale", None)
    if scale is None:
        scale = inner.sm_scale
    sinks = getattr(inner, "sinks", None)
    if sinks is not None and sinks.dtype != torch.float32:
        sinks = sinks.float()

    cache_block_size = kv_cache.block_size_for_layer(layer_idx)
    if meta is None:
        group_index = kv_cache.group_index_for_layer(layer_idx)
        if ctx.kv_groups is None:
            raw_slots, raw_tables = ctx.slot_mapping, ctx.block_tables
        else:
            group = ctx.kv_groups[group_index]
            raw_slots, raw_tables, cache_block_size = group.slot_mapping, group.block_tables, group.block_size
        meta = _kernel_metadata(ctx, None if ctx.kv_groups is None else group_index, raw_slots,
                                raw_tables, cache_block_size, device=x.device)

    if hasattr(inner, "qkv_proj"):
        qkv = inner.qkv_proj(x2)
        hd = qkv.shape[-1] // (n_heads + 2 * n_kv_heads)
        q, k, v = qkv.split([n_heads * hd, n_kv_heads * hd, n_kv_heads * hd], dim=-1)
    else:
        q, k, v = inner.q_proj(x2), inner.k_proj(x2), inner.v_proj(x2)
        hd = q.shape[-1] // n_heads
    gate = None
    if getattr(inner, "attn_output_gate", False):
        # Qwen3-Next style: q_proj emits (query, gate) per head; out *= sigmoid(gate) (sdpa.py:735-736)
        hd = q.shape[-1] // (2 * n_heads)
        q, gate = q.reshape(t, n_heads, 2 * hd).split([hd, hd], dim=-1)
        q, gate = q.reshape(t, n_heads * hd), gate.reshape(t, n_heads * hd)
    q = q.reshape(t, n_heads, hd)
    k = k.reshape(t, n_kv_heads, hd)
    v = v.reshape(t, n_kv_heads, hd)
    rope = getattr(inner, "rope", None)
    q_norm, k_norm = getattr(inner, "q_norm", None), getattr(inner, "k_norm", None)
    k_cache, v_cache = kv_cache.key_caches[layer_idx], kv_cache.value_caches[layer_idx]
    cache_hd = kv_cache.head_dim_per_layer[layer_idx]
    if hd > cache_hd:
        raise ValueError(f"layer head_dim {hd} exceeds the cache head_dim {cache_hd}")
    fused = (
        _FUSE_NORM_ROPE_CACHE
        and hd == cache_hd
        and q.dtype == kv_cache.dtype
        and all(x.stride(2) == 1 and x.stride(1) == hd for x in (q, k, v))
        and (q_norm is None) == (k_norm is None)
        and (q_norm is None or (hasattr(q_norm, "weight") and hasattr(q_norm, "eps")))
    )
    if fused:
        # one launch: per-head norm -> RoPE -> cache write, q/k/v read as strided views of qkv
        k_cache, v_cache = ops.norm_rope_cache(
            q, k, v, k_cache, v_cache, meta.slot_mapping, positions=meta.positions,
            rot_dim=0 if rope is None else rope.dims, base=10000.0 if rope is None else rope.base,
            traditional=False if rope is None else rope.traditional,
            q_norm_weight=None if q_norm is None else q_norm.weight,
            k_norm_weight=None if k_norm is None else k_norm.weight,
            eps=1e-6 if q_norm is None else q_norm.eps)
    else:
        if q_norm is not None:  # Qwen3 per-head norms (sdpa.py:348-353)
            q = q_norm(q)
            k = k_norm(k)
        if rope is not None:
            if not q.is_contiguous():
                q = q.contiguous()
            if not k.is_contiguous():
                k = k.contiguous()
            ops.rope(q, k, meta.positions, rope.dims, rope.base, rope.traditional)
        if q.dtype != kv_cache.dtype:
            q, k, v = q.to(kv_cache.dtype), k.to(kv_cache.dtype), v.to(kv_cache.dtype)
        if hd < cache_hd:
            # variable-head_dim models allocate the cache at the widest layer: zero-pad q/k/v up to it; zeros change
            # neither q.k nor the first hd columns of P.V (sdpa.py:377-420, :536-539)
            pad = (0, cache_hd - hd)
            q, k, v = (torch.nn.functional.pad(z, pad) for z in (q, k, v))
        k_cache, v_cache = ops.reshape_and_cache(k.contiguous(), v.contiguous(), k_cache, v_cache, meta.slot_mapping)
    kv_cache.replace_layer_cache(layer_idx, k_cache, v_cache)

    if meta.block_size != cache_block_size:  # zero-copy view at the kernel page size
        kvh, d = k_cache.shape[2], k_cache.shape[3]
        k_cache = k_cache.reshape(-1, meta.block_size, kvh, d)
        v_cache = v_cache.reshape(-1, meta.block_size, kvh, d)

    # fp16 output straight from the attention kernels when o_proj runs a tensor-core kernel (fp16 operands) and
    # nothing sits in between: the same bf16-rounded values, no cast launch
    o_proj = inner.o_proj
    plain_tail = hd == cache_hd and gate is None and not (getattr(inner, "gating", False) and hasattr(inner, "g_proj"))
    f16_out = plain_tail and q.dtype == torch.bfloat16 and (
        getattr(o_proj, "weight_blob", None) is not None or getattr(o_proj, "weight_tc", None) is not None)
    out = torch.empty((t, n_heads, cache_hd), dtype=torch.float16 if f16_out else q.dtype, device=q.device)
    ops.paged_attention_primitive(
        q, k_cache, v_cache, kv_cache.kv_heads_per_layer[layer_idx], float(scale), 0.0,
        meta.block_tables, meta.seq_lens, meta.cu_seqlens_q, meta.block_size, meta.max_seq_len,
        kv_cache.sliding_window_per_layer[layer_idx], out,
        window_seqlen_q=getattr(ctx, "verify_window_q", 1), sinks=sinks)  # sdpa.py:713,730
    if hd < cache_hd:
        out = out[..., :hd]  # strip the pad before o_proj (truncate_padded_output)
    out = out.reshape(t, n_heads * hd)
    if gate is not None:
        out = out * torch.sigmoid(gate.to(out.dtype))
    if getattr(inner, "gating", False) and hasattr(inner, "g_proj"):
        # Laguna-style per-head gate: softplus(g_proj(x)) in fp32, one scalar per head (sdpa.py:742-773)
        g = torch.nn.functional.softplus(inner.g_proj(x2).float()).to(out.dtype)
        out = (out.reshape(t, n_heads, hd) * g.unsqueeze(-1)).reshape(t, n_heads * hd)
    y = inner.o_proj(out.contiguous())
    return y.reshape(*lead_shape, y.shape[-1])
