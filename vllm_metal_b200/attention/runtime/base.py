"""Shared behaviour of the paged-attention runtimes (reference: ``PagedAttentionRuntimeBase``,
vllm_metal/attention/runtime/base.py:14-96)."""

from __future__ import annotations

from collections.abc import Sequence
from typing import Any


class PagedAttentionRuntimeBase:
    _cache: Any = None
    _block_size: int = 16

    def _require_initialized(self, what: str) -> Any:
        if self._cache is None:
            raise RuntimeError(f"{what} called before initialize()")
        return self._cache

    def num_blocks(self) -> int:
        return self._require_initialized("num_blocks").num_blocks

    def kv_scheduler_group_indices(self) -> tuple[int, ...]:
        return (0,)  # one KV group unless a grouped layout was adopted

    def kv_group_block_sizes(self) -> tuple[int, ...]:
        return (self._block_size,)

    def copy_blocks(self, block_copies: Sequence[tuple[int, int]]) -> None:
        self._require_initialized("copy_blocks").copy_blocks(block_copies)

    def needs_step_context(self) -> bool:
        return False  # plain MHA keeps no per-request state besides the paged cache

    def release_requests(self, req_ids: set[str]) -> None:
        return None

    def warm_up(self) -> None:
        """Load the CUDA library and touch the kernels' lazy state before the first request."""
        from ...metal import warm_up_kernels

        warm_up_kernels()
