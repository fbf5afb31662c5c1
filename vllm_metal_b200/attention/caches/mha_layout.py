"""vLLM's ``KVCacheConfig`` for standard (full + sliding-window) MHA, translated into the per-layer facts the
paged cache needs (reference: ``MHAKVCacheLayout``, vllm_metal/attention/caches/mha_layout.py:57-208).

vLLM owns grouping and capacity planning: layers of one ``kv_cache_group`` share a spec (and the scheduler
hands out one block-id list per group), and every ``KVCacheTensor`` names the layers stored in it.  Nothing is
regrouped here -- the translation only checks that the config describes exactly the model's attention layers
and records, per layer, which group's slot mapping / block tables it reads (``group_index``), its page size,
head geometry and window.
"""

from __future__ import annotations

from dataclasses import dataclass

from vllm.v1.kv_cache_interface import FullAttentionSpec, KVCacheConfig, SlidingWindowSpec

NO_SLIDING_WINDOW = -1


@dataclass(frozen=True)
class MHALayerKVLayout:
    tensor_index: int
    group_index: int
    block_size: int
    num_kv_heads: int
    head_dim: int
    sliding_window: int

    def cache_shape(self, num_blocks: int) -> tuple[int, int, int, int]:
        return (num_blocks, self.block_size, self.num_kv_heads, self.head_dim)


@dataclass(frozen=True)
class MHAKVCacheLayout:
    num_blocks: int
    tensor_sizes: tuple[int, ...]
    layers: tuple[MHALayerKVLayout, ...]
    group_block_sizes: tuple[int, ...]
    slot_layers: tuple[tuple[int, ...], ...]

    @property
    def total_bytes(self) -> int:
        return sum(self.tensor_sizes)

    @classmethod
    def from_config(cls, config: KVCacheConfig, model_layer_names: tuple[str, ...]) -> MHAKVCacheLayout:
        order = {name: i for i, name in enumerate(model_layer_names)}
        specs, group_of = [], {}
        for gi, group in enumerate(config.kv_cache_groups):
            spec = group.kv_cache_spec
            if not isinstance(spec, (FullAttentionSpec, SlidingWindowSpec)):
                raise NotImplementedError("standard MHA layout requires FullAttentionSpec or SlidingWindowSpec groups")
            if getattr(spec, "head_size_v", spec.head_size) != spec.head_size:
                raise NotImplementedError("standard MHA layout requires matching key and value head sizes")
            specs.append(spec)
            for name in group.layer_names:
                group_of[name] = gi
        if set(group_of) != set(order):
            raise ValueError("group layer mapping must contain the same layers as model_layer_names")
        tensor_of, slot_layers = {}, []
        for ti, tensor in enumerate(config.kv_cache_tensors):
            if getattr(tensor, "offset", 0) or getattr(tensor, "block_stride", 0):
                raise NotImplementedError("standard MHA layout does not support KV tensors with offset and block_stride")
            members = []
            for name in tensor.shared_by:
                spec = specs[group_of[name]]
                want = config.num_blocks * spec.page_size_bytes
                if tensor.size != want:
                    raise ValueError(f"KV cache tensor {ti} size {tensor.size} does not match {config.num_blocks} blocks "
                                     f"of {spec.page_size_bytes} bytes for layer {name!r}")
                tensor_of[name] = ti
                members.append(order[name])
            slot_layers.append(tuple(members))
        if set(tensor_of) != set(order):
            raise ValueError("tensor layer mapping must contain the same layers as model_layer_names")
        layers = []
        for name in model_layer_names:
            spec = specs[group_of[name]]
            layers.append(MHALayerKVLayout(
                tensor_index=tensor_of[name], group_index=group_of[name], block_size=spec.block_size,
                num_kv_heads=spec.num_kv_heads, head_dim=spec.head_size,
                sliding_window=spec.sliding_window if isinstance(spec, SlidingWindowSpec) else NO_SLIDING_WINDOW))
        return cls(num_blocks=config.num_blocks, tensor_sizes=tuple(t.size for t in config.kv_cache_tensors),
                   layers=tuple(layers), group_block_sizes=tuple(s.block_size for s in specs),
                   slot_layers=tuple(slot_layers))
