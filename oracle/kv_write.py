"""CPU oracle: ``reshape_and_cache`` KV scatter.  TEST INFRASTRUCTURE ONLY.

Restates ``vllm_metal/metal/kernels_v2/reshape_and_cache.metal:36-96`` and the
scatter it must equal bit for bit (``tests/test_reshape_and_cache.py:15-27``):
for each token ``t``: ``slot = slot_mapping[t]``; negative slots are padding and
skipped (``reshape_and_cache.metal:61-65``); otherwise row ``t`` of ``key`` /
``value`` (``[T, H_kv, D]``) is copied to ``cache[slot // bs, slot % bs]``.
Works on any NumPy dtype (use uint16 views for bf16) so equality is bitwise.
"""

from __future__ import annotations

import numpy as np


def reshape_and_cache(
    key: np.ndarray,
    value: np.ndarray,
    key_cache: np.ndarray,
    value_cache: np.ndarray,
    slot_mapping: np.ndarray,
) -> tuple[np.ndarray, np.ndarray]:
    """In-place scatter; returns the (aliased) caches like the reference op."""
    num_blocks, block_size, h, d = key_cache.shape
    assert key.shape[1:] == (h, d) and value.shape[1:] == (h, d)
    for t in range(key.shape[0]):
        slot = int(slot_mapping[t])
        if slot < 0:
            continue
        b, o = divmod(slot, block_size)
        key_cache[b, o] = key[t]
        value_cache[b, o] = value[t]
    return key_cache, value_cache


def copy_blocks(key_cache: np.ndarray, value_cache: np.ndarray, src: list[int], dst: list[int]) -> None:
    """Copy-on-write block copy, ``arr[dst] = arr[src]`` (kv_cache.py:327-354)."""
    if not src:
        return
    s = np.asarray(src, dtype=np.int64)
    d = np.asarray(dst, dtype=np.int64)
    key_cache[d] = key_cache[s]
    value_cache[d] = value_cache[s]
