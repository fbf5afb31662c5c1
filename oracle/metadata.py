"""CPU oracle: per-step paged metadata.  TEST INFRASTRUCTURE ONLY.

Restates, as plain functions returning plain lists:

* ``prepare_grouped`` / ``prepare_unified`` -- ``vllm_metal/attention/context.py:162-314``
  (decode rows first, then prefill; ``slot = block_ids[pos // bs] * bs + pos % bs``;
  ``context_lens`` include this step's tokens; RoPE ``offsets``).
* ``pick_kernel_block_size`` / ``build_block_tables`` --
  ``vllm_metal/attention/impls/sdpa.py:115-163`` (pad rows with 0; when the cache
  block is not one of {32,16,8} pick the largest divisor and expand block ``b``
  into ``b*r .. b*r+r-1``).

Pinned by the reference's literal tests (``tests/test_paged_attention.py:132-323``,
``tests/test_block_size_translation.py:30-108``) and by fixtures generated from
the reference's own functions (``tests/golden/make_golden.py``).
"""

from __future__ import annotations

import numpy as np

KERNEL_BLOCK_SIZES = (32, 16, 8)  # sdpa.py:49


def prepare_grouped(decode_requests, prefill_requests, block_sizes, merge_verify_windows=False):
    """Returns a dict with the fields of ``PagedAttentionContext`` (context.py:49-100)."""
    slot_maps = [[] for _ in block_sizes]
    tables = [[] for _ in block_sizes]
    cu = [0]
    ctx_lens = []
    offsets = []
    max_window = 1
    for req in decode_requests:
        if len(req) == 2:
            ids_by_group, seq_len = req
            n = 1
        else:
            ids_by_group, seq_len, n = req
        for gi, bs in enumerate(block_sizes):
            ids = ids_by_group[gi]
            for pos in range(seq_len, seq_len + n):
                slot_maps[gi].append(ids[pos // bs] * bs + pos % bs)
            if n > 1 and not merge_verify_windows:
                tables[gi].extend(list(ids) for _ in range(n))
            else:
                tables[gi].append(list(ids))
        if n > 1 and not merge_verify_windows:
            for pos in range(seq_len, seq_len + n):
                cu.append(cu[-1] + 1)
                ctx_lens.append(pos + 1)
                offsets.append(pos)
            continue
        cu.append(cu[-1] + n)
        ctx_lens.append(seq_len + n)
        offsets.append(seq_len)
        max_window = max(max_window, n)
    for ids_by_group, n, start in prefill_requests:
        for gi, bs in enumerate(block_sizes):
            ids = ids_by_group[gi]
            for pos in range(start, start + n):
                slot_maps[gi].append(ids[pos // bs] * bs + pos % bs)
            tables[gi].append(list(ids))
        cu.append(cu[-1] + n)
        ctx_lens.append(start + n)
        offsets.append(start)
    groups = [
        {"slot_mapping": slot_maps[gi], "block_tables": tables[gi], "block_size": bs}
        for gi, bs in enumerate(block_sizes)
    ]
    return {
        "slot_mapping": groups[0]["slot_mapping"],
        "block_tables": groups[0]["block_tables"],
        "context_lens": ctx_lens,
        "cu_seqlens": cu,
        "offsets": offsets,
        "num_decode_requests": len(decode_requests),
        "verify_window_q": 1 if prefill_requests else max_window,
        "kv_groups": groups,
    }


def prepare_unified(decode_requests, prefill_requests, block_size, merge_verify_windows=False):
    """Single-group wrapper (context.py:286-314)."""
    dec = []
    for r in decode_requests:
        if len(r) == 2:
            dec.append(([r[0]], r[1]))
        else:
            dec.append(([r[0]], r[1], r[2]))
    pre = [([ids], n, start) for ids, n, start in prefill_requests]
    return prepare_grouped(dec, pre, (block_size,), merge_verify_windows)


def pick_kernel_block_size(cache_block_size: int) -> int:
    """sdpa.py:115-124."""
    for kbs in KERNEL_BLOCK_SIZES:
        if cache_block_size % kbs == 0:
            return kbs
    raise ValueError(
        f"Cache block_size={cache_block_size} is not divisible by any "
        f"supported kernel block size {KERNEL_BLOCK_SIZES}."
    )


def build_block_tables(raw_block_tables, cache_block_size):
    """sdpa.py:127-163 -> (int32 ndarray [S, max_blocks(*ratio)], kernel_block_size)."""
    if not raw_block_tables:
        return np.zeros((0, 0), dtype=np.int32), cache_block_size
    max_blocks = max(len(bt) for bt in raw_block_tables)
    padded = np.array(
        [list(bt) + [0] * (max_blocks - len(bt)) for bt in raw_block_tables], dtype=np.int32
    ).reshape(len(raw_block_tables), max_blocks)
    if cache_block_size in KERNEL_BLOCK_SIZES:
        return padded, cache_block_size
    kbs = pick_kernel_block_size(cache_block_size)
    ratio = cache_block_size // kbs
    expanded = (padded[:, :, None] * ratio + np.arange(ratio, dtype=np.int32)[None, None, :]).reshape(
        padded.shape[0], -1
    )
    return expanded.astype(np.int32), kbs
