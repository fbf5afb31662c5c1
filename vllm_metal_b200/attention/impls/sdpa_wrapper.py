"""Attention-module wrapper that routes a model's ``self_attn`` call through the paged CUDA path (reference:
``SDPAPagedAttentionWrapper`` / ``patch_sdpa_attention``, vllm_metal/attention/impls/sdpa_wrapper.py:45-225).

With a ``PagedAttentionContext`` set for the step (``attention/context.py``) the call runs ``sdpa_forward``:
quantized projections -> per-head norms -> RoPE -> ``reshape_and_cache`` -> ``paged_attention_primitive`` ->
``o_proj``.  Without one it delegates to the wrapped module unchanged, so code that never enters a paged step
(a dense warm-up, an encoder pass) keeps working -- the reference's contract.
"""

from __future__ import annotations

from typing import Any

import torch
from torch import nn

from ..caches.kv_cache import B200PagedKVCache
from ..context import get_context
from ..patching import walk_and_wrap
from .sdpa import sdpa_forward


class SDPAPagedAttentionWrapper(nn.Module):
    def __init__(self, inner: nn.Module, layer_idx: int, kv_cache: B200PagedKVCache, block_size: int, *,
                 cache_idx: int | None = None) -> None:
        super().__init__()
        self._inner = inner
        # plain attributes (not modules/buffers): the cache is shared by every wrapper and rebound on re-init
        object.__setattr__(self, "_mk_layer_idx", layer_idx)
        object.__setattr__(self, "_mk_kv_cache", kv_cache)
        object.__setattr__(self, "_mk_block_size", block_size)
        object.__setattr__(self, "_mk_cache_idx", cache_idx if cache_idx is not None else layer_idx)

    # what host model code reads off the attention module (reference :84-104)
    @property
    def rope(self) -> Any:
        return self._inner.rope

    @property
    def rotary_emb(self) -> Any:
        return self._inner.rotary_emb

    @property
    def is_local(self) -> Any:
        return getattr(self._inner, "is_local", False)

    def rebind_cache(self, kv_cache: B200PagedKVCache, block_size: int, *, cache_idx: int | None = None) -> None:
        object.__setattr__(self, "_mk_kv_cache", kv_cache)
        object.__setattr__(self, "_mk_block_size", block_size)
        object.__setattr__(self, "_mk_cache_idx", cache_idx if cache_idx is not None else self._mk_layer_idx)

    def forward(self, x: torch.Tensor, mask: Any = None, cache: Any = None, **kwargs: Any) -> Any:
        ctx = get_context()
        if ctx is None:  # no paged step in flight: the original attention
            return self._inner(x, mask=mask, cache=cache, **kwargs)
        return sdpa_forward(self._inner, x, ctx, self._mk_kv_cache, self._mk_cache_idx)


def patch_sdpa_attention(model: Any, kv_cache: B200PagedKVCache, block_size: int, *,
                         cache_idx_map: dict[int, int] | None = None, only_layers: list[int] | None = None) -> int:
    """Replace every attention module with a wrapper (or rebind an existing wrapper's cache in place).
    ``cache_idx_map``: model layer index -> index into a compact cache (hybrid models)."""

    def wrap_layer(layer_idx: int, attn: Any) -> Any:
        cache_idx = cache_idx_map[layer_idx] if cache_idx_map is not None and layer_idx in cache_idx_map else layer_idx
        if isinstance(attn, SDPAPagedAttentionWrapper):
            attn.rebind_cache(kv_cache, block_size, cache_idx=cache_idx)
            return attn
        return SDPAPagedAttentionWrapper(attn, layer_idx, kv_cache, block_size, cache_idx=cache_idx)

    return walk_and_wrap(model, wrap_layer, only_layers=only_layers)
