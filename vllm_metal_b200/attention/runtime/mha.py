"""Runtime for standard MHA / GQA models: allocates the paged cache, patches the model's attention modules onto
the CUDA path and warms the kernels up (reference: ``MHAPagedAttentionRuntime``,
vllm_metal/attention/runtime/mha.py:16-105)."""

from __future__ import annotations

from typing import Any

import torch

from ..caches.kv_cache import B200PagedKVCache
from ..caches.mha_layout import MHAKVCacheLayout
from ..impls.sdpa_wrapper import patch_sdpa_attention
from .base import PagedAttentionRuntimeBase


class MHAPagedAttentionRuntime(PagedAttentionRuntimeBase):
    def __init__(self, *, num_layers: int, num_kv_heads: int, head_dim: int, block_size: int,
                 dtype: torch.dtype = torch.bfloat16, device: torch.device | str = "cuda",
                 cache_idx_map: dict[int, int] | None = None, kv_heads_per_layer: list[int] | None = None,
                 head_dim_per_layer: list[int] | None = None, sliding_window_per_layer: list[int] | None = None) -> None:
        self._num_layers, self._num_kv_heads, self._head_dim = num_layers, num_kv_heads, head_dim
        self._block_size, self._dtype, self._device = block_size, dtype, device
        self._cache_idx_map = cache_idx_map
        self._kv_heads_per_layer, self._head_dim_per_layer = kv_heads_per_layer, head_dim_per_layer
        self._sliding_window_per_layer = sliding_window_per_layer
        self._cache: B200PagedKVCache | None = None
        self._layout: MHAKVCacheLayout | None = None

    def initialize(self, num_blocks: int) -> None:
        self._cache = B200PagedKVCache(
            self._num_layers, self._num_kv_heads, self._head_dim, num_blocks, self._block_size, self._dtype, self._device,
            kv_heads_per_layer=self._kv_heads_per_layer, head_dim_per_layer=self._head_dim_per_layer,
            sliding_window_per_layer=self._sliding_window_per_layer)

    def adopt_layout(self, layout: MHAKVCacheLayout) -> None:
        """Use vLLM's grouped MHA layout (mixed full / sliding-window groups) as the runtime cache."""
        self._layout = layout
        self._block_size = layout.group_block_sizes[0]
        self._cache = B200PagedKVCache.from_layout(layout, self._dtype, self._device)

    def kv_scheduler_group_indices(self) -> tuple[int, ...]:
        if self._layout is None:
            return super().kv_scheduler_group_indices()
        self._require_initialized("kv_scheduler_group_indices")
        return tuple(range(len(self._layout.group_block_sizes)))

    def kv_group_block_sizes(self) -> tuple[int, ...]:
        if self._layout is None:
            return super().kv_group_block_sizes()
        self._require_initialized("kv_group_block_sizes")
        return self._layout.group_block_sizes

    def patch_model(self, model: Any) -> int:
        cache = self._require_initialized("patch_model")
        return patch_sdpa_attention(model, cache, self._block_size, cache_idx_map=self._cache_idx_map)

    @property
    def kv_cache(self) -> B200PagedKVCache:
        return self._require_initialized("kv_cache")
