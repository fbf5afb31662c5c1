"""Paged KV cache in HBM (reference: ``MetalPagedKVCache``,
vllm_metal/attention/caches/kv_cache.py:37-373).

Per layer one K and one V tensor ``[num_blocks, block_size, H_kv, D]``
(token-major rows, kv_cache.py:223-234) in the model dtype.  On B200 all layers
are carved out of ONE allocation per tensor kind (``[L, num_blocks, ...]``) so a
180 GB part is sized with a single ``cudaMalloc`` and CoW block copies are one
launch per layer; per-layer tensors are views into it.  Per-layer
``kv_heads`` / ``head_dim`` / ``sliding_window`` lists are kept for models with
mixed full/sliding layers (gpt-oss).
"""

from __future__ import annotations

from collections.abc import Sequence

import torch

from ...metal import get_ops


class B200PagedKVCache:
    def __init__(
        self,
        num_layers: int,
        num_kv_heads: int,
        head_dim: int,
        num_blocks: int,
        block_size: int,
        dtype: torch.dtype = torch.bfloat16,
        device: torch.device | str = "cuda",
        *,
        kv_heads_per_layer: list[int] | None = None,
        head_dim_per_layer: list[int] | None = None,
        sliding_window_per_layer: list[int] | None = None,
        block_size_per_layer: list[int] | None = None,
    ) -> None:
        if dtype not in (torch.float16, torch.bfloat16, torch.float32):
            raise ValueError(f"unsupported KV cache dtype {dtype}")
        self.num_layers = num_layers
        self.num_kv_heads = num_kv_heads
        self.head_dim = head_dim
        self.num_blocks = num_blocks
        self.block_size = block_size
        self.dtype = dtype
        self.device = torch.device(device)
        self.kv_heads_per_layer = kv_heads_per_layer or [num_kv_heads] * num_layers
        self.head_dim_per_layer = head_dim_per_layer or [head_dim] * num_layers
        self.sliding_window_per_layer = sliding_window_per_layer or [-1] * num_layers
        self.block_size_per_layer = block_size_per_layer or [block_size] * num_layers
        self._group_index_per_layer = [0] * num_layers
        for name, lst in (
            ("kv_heads_per_layer", self.kv_heads_per_layer),
            ("head_dim_per_layer", self.head_dim_per_layer),
            ("sliding_window_per_layer", self.sliding_window_per_layer),
            ("block_size_per_layer", self.block_size_per_layer),
        ):
            if len(lst) != num_layers:
                raise ValueError(f"{name} length {len(lst)} != num_layers {num_layers}")

        self.key_caches: list[torch.Tensor] = []
        self.value_caches: list[torch.Tensor] = []
        uniform = len(set(self.kv_heads_per_layer)) == 1 and len(set(self.head_dim_per_layer)) == 1 and \
            len(set(self.block_size_per_layer)) == 1
        if uniform:
            shape = (num_layers, num_blocks, block_size, self.kv_heads_per_layer[0], self.head_dim_per_layer[0])
            self._k_all = torch.zeros(shape, dtype=dtype, device=self.device)
            self._v_all = torch.zeros(shape, dtype=dtype, device=self.device)
            self.key_caches = [self._k_all[i] for i in range(num_layers)]
            self.value_caches = [self._v_all[i] for i in range(num_layers)]
        else:
            for i in range(num_layers):
                shape = (num_blocks, self.block_size_per_layer[i], self.kv_heads_per_layer[i], self.head_dim_per_layer[i])
                self.key_caches.append(torch.zeros(shape, dtype=dtype, device=self.device))
                self.value_caches.append(torch.zeros(shape, dtype=dtype, device=self.device))

    @classmethod
    def from_layout(cls, layout, dtype: torch.dtype = torch.bfloat16, device: torch.device | str = "cuda"):
        """Cache shaped by vLLM's grouped layout (reference ``MetalPagedKVCache.from_layout``,
        kv_cache.py:128-170): per-layer page size / heads / window and the scheduler KV group each layer reads
        its slot mapping and block tables from (mixed full + sliding-window models)."""
        ls = layout.layers
        self = cls(len(ls), ls[0].num_kv_heads, ls[0].head_dim, layout.num_blocks, ls[0].block_size, dtype, device,
                   kv_heads_per_layer=[l.num_kv_heads for l in ls], head_dim_per_layer=[l.head_dim for l in ls],
                   sliding_window_per_layer=[l.sliding_window for l in ls],
                   block_size_per_layer=[l.block_size for l in ls])
        self._group_index_per_layer = [l.group_index for l in ls]
        return self

    # -- reference API ----------------------------------------------------------
    def block_size_for_layer(self, layer_idx: int) -> int:
        return self.block_size_per_layer[layer_idx]

    def group_index_for_layer(self, layer_idx: int) -> int:
        return self._group_index_per_layer[layer_idx]

    def replace_layer_cache(self, layer_idx: int, key_cache: torch.Tensor, value_cache: torch.Tensor) -> None:
        """Rebind the handles ``reshape_and_cache`` returns (kv_cache.py:310-325).
        The op writes in place, so these alias the tensors already held."""
        self.key_caches[layer_idx] = key_cache
        self.value_caches[layer_idx] = value_cache

    def copy_blocks(self, block_copies: Sequence[tuple[int, int]]) -> None:
        """Scheduler copy-on-write: ``cache[dst] = cache[src]`` (kv_cache.py:327-354)."""
        if not block_copies:
            return
        src_ids, dst_ids = zip(*block_copies, strict=True)
        if any(b < 0 or b >= self.num_blocks for b in (*src_ids, *dst_ids)):
            raise RuntimeError("paged KV block copy contains an out-of-range block id")
        if len(set(dst_ids)) != len(dst_ids):
            raise RuntimeError("paged KV block copy writes one destination block twice")
        # The reference's ``array[dst] = array[src]`` gathers every source before it scatters; the kernel copies
        # all pairs in parallel, so a pair whose source is another pair's destination must still read the OLD
        # contents: those sources are staged first.
        dst_set = set(dst_ids)
        ops = get_ops()
        clash = [(s_, d_) for s_, d_ in block_copies if s_ in dst_set and s_ != d_]
        if clash:
            # gather-then-scatter through a scratch copy of the clashing sources
            cs = sorted({s_ for s_, _ in clash})
            idx = torch.tensor(cs, dtype=torch.int64, device=self.device)
            pos = {b: i for i, b in enumerate(cs)}
            for k, v in zip(self.key_caches, self.value_caches, strict=True):
                ks, vs = k.index_select(0, idx), v.index_select(0, idx)
                plain = [(s_, d_) for s_, d_ in block_copies if s_ not in dst_set and s_ != d_]
                if plain:
                    ps = torch.tensor([s_ for s_, _ in plain], dtype=torch.int32, device=self.device)
                    pd = torch.tensor([d_ for _, d_ in plain], dtype=torch.int32, device=self.device)
                    ops.copy_blocks(k, v, ps, pd)
                di = torch.tensor([d_ for _, d_ in clash], dtype=torch.int64, device=self.device)
                si = torch.tensor([pos[s_] for s_, _ in clash], dtype=torch.int64, device=self.device)
                k.index_copy_(0, di, ks.index_select(0, si))
                v.index_copy_(0, di, vs.index_select(0, si))
            return
        src = torch.tensor(src_ids, dtype=torch.int32, device=self.device)
        dst = torch.tensor(dst_ids, dtype=torch.int32, device=self.device)
        for k, v in zip(self.key_caches, self.value_caches, strict=True):
            ops.copy_blocks(k, v, src, dst)

    def bytes_per_block(self) -> int:
        esize = torch.empty((), dtype=self.dtype).element_size()
        return sum(
            2 * b * h * d * esize
            for b, h, d in zip(self.block_size_per_layer, self.kv_heads_per_layer, self.head_dim_per_layer, strict=True)
        )

    @property
    def nbytes(self) -> int:
        return self.bytes_per_block() * self.num_blocks
