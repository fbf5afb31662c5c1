"""Per-step paged-attention context (host side).

Same contract as the reference's ``vllm_metal/attention/context.py``
(``PagedAttentionContext`` :49-100, ``prepare_grouped`` :162-283,
``prepare_unified`` :286-314): before a forward pass the runner packs decode
rows first, then prefill chunks, and publishes ``slot_mapping`` /
``block_tables`` / ``cu_seqlens`` / ``context_lens`` / RoPE ``offsets`` through
a thread-local that the attention wrappers read.

B200 differences: the per-token work is vectorised with NumPy (a 2048-token
chunk is one ``arange``), and the context memoises the DEVICE tensors each KV
group needs (``kernel_metadata_cache``), so a 32-80 layer model uploads its
metadata once per step from pinned memory instead of once per layer.
"""

from __future__ import annotations

import threading
from collections.abc import Sequence
from dataclasses import dataclass, field
from typing import Any

import numpy as np

_tls = threading.local()


@dataclass
class PagedKVGroupContext:
    """Paged metadata of one scheduler KV-cache group for one packed forward."""

    slot_mapping: list[int]
    block_tables: list[list[int]]
    block_size: int


@dataclass
class PagedAttentionContext:
    """What the attention wrappers read during one forward pass."""

    slot_mapping: list[int]
    block_tables: list[list[int]] = field(default_factory=list)
    # KV length of each segment INCLUDING this step's tokens.
    context_lens: list[int] = field(default_factory=list)
    # RoPE position of each segment's first token.
    offsets: list[int] = field(default_factory=list)
    cu_seqlens: list[int] | None = None
    num_decode_requests: int = 0
    verify_window_q: int = 1
    kv_groups: tuple[PagedKVGroupContext, ...] | None = None
    # (group index, cache block size) -> device tensors, filled by impls.sdpa
    kernel_metadata_cache: dict[tuple[int | None, int], Any] = field(default_factory=dict)
    # per-token RoPE positions (device int32), memoised by impls.sdpa
    positions_cache: Any = None


def set_context(ctx: PagedAttentionContext) -> None:
    _tls.ctx = ctx


def get_context() -> PagedAttentionContext | None:
    return getattr(_tls, "ctx", None)


def clear_context() -> None:
    _tls.ctx = None


def _slots(block_ids: Sequence[int], first: int, count: int, block_size: int) -> list[int]:
    """Flat cache slots of positions ``first .. first+count-1`` of one request."""
    pos = np.arange(first, first + count, dtype=np.int64)
    ids = np.asarray(block_ids, dtype=np.int64)
    return (ids[pos // block_size] * block_size + pos % block_size).tolist()


def prepare_grouped(
    decode_requests: Sequence[tuple],
    prefill_requests: Sequence[tuple],
    block_sizes: Sequence[int],
    merge_verify_windows: bool = False,
) -> None:
    """Stage the context for one forward pass over every KV-cache group.

    ``decode_requests``: ``(block_ids_by_group, seq_len[, num_tokens])`` with
    ``seq_len`` = tokens already cached; ``num_tokens > 1`` is a speculative
    verification window (expanded to one-row segments unless
    ``merge_verify_windows``).  ``prefill_requests``: ``(block_ids_by_group,
    num_tokens, start_pos)``.  Semantics follow context.py:198-283.
    """
    n_groups = len(block_sizes)
    slot_maps: list[list[int]] = [[] for _ in range(n_groups)]
    tables: list[list[list[int]]] = [[] for _ in range(n_groups)]
    cu = [0]
    ctx_lens: list[int] = []
    offsets: list[int] = []
    widest_window = 1

    for req in decode_requests:
        ids_by_group, seq_len = req[0], req[1]
        count = req[2] if len(req) > 2 else 1
        expand = count > 1 and not merge_verify_windows
        for gi, bs in enumerate(block_sizes):
            slot_maps[gi].extend(_slots(ids_by_group[gi], seq_len, count, bs))
            row = list(ids_by_group[gi])
            tables[gi].extend([row] * (count if expand else 1))
        if expand:
            for pos in range(seq_len, seq_len + count):
                cu.append(cu[-1] + 1)
                ctx_lens.append(pos + 1)
                offsets.append(pos)
        else:
            cu.append(cu[-1] + count)
            ctx_lens.append(seq_len + count)
            offsets.append(seq_len)
            widest_window = max(widest_window, count)

    for ids_by_group, count, start in prefill_requests:
        for gi, bs in enumerate(block_sizes):
            slot_maps[gi].extend(_slots(ids_by_group[gi], start, count, bs))
            tables[gi].append(list(ids_by_group[gi]))
        cu.append(cu[-1] + count)
        ctx_lens.append(start + count)
        offsets.append(start)

    groups = tuple(
        PagedKVGroupContext(slot_mapping=slot_maps[gi], block_tables=tables[gi], block_size=bs)
        for gi, bs in enumerate(block_sizes)
    )
    set_context(
        PagedAttentionContext(
            slot_mapping=groups[0].slot_mapping,
            block_tables=groups[0].block_tables,
            context_lens=ctx_lens,
            cu_seqlens=cu,
            offsets=offsets,
            num_decode_requests=len(decode_requests),
            # any prefill segment keeps the batch off the verification-window route
            verify_window_q=1 if prefill_requests else widest_window,
            kv_groups=groups,
        )
    )


def prepare_unified(
    decode_requests: Sequence[tuple],
    prefill_requests: Sequence[tuple],
    block_size: int,
    *,
    merge_verify_windows: bool = False,
) -> None:
    """Single KV-cache group convenience wrapper (context.py:286-314)."""
    dec = [([r[0]], *r[1:]) for r in decode_requests]
    pre = [([ids], count, start) for ids, count, start in prefill_requests]
    prepare_grouped(dec, pre, (block_size,), merge_verify_windows=merge_verify_windows)
