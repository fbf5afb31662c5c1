"""Host-side attention layer on CPU with a spy in place of the CUDA ops: per-group slot / page-table routing
(the reference's literals, tests/test_attention_sdpa.py:630-685), layout translation from vLLM's KVCacheConfig
(attention/caches/mha_layout.py), wrapper / patching / runtime contract (sdpa_wrapper.py:45-225, patching.py:57,
runtime/mha.py:16-98), head-dim padding and output gates (sdpa.py:377-420, :735-773).  No kernels run."""

from __future__ import annotations

from types import SimpleNamespace
from unittest.mock import patch

import pytest
import torch
from torch import nn

pytest.importorskip("vllm")

from vllm_metal_b200.attention import context as ctx_mod  # noqa: E402
from vllm_metal_b200.attention.caches.kv_cache import B200PagedKVCache  # noqa: E402
from vllm_metal_b200.attention.caches.mha_layout import MHAKVCacheLayout, MHALayerKVLayout  # noqa: E402
from vllm_metal_b200.attention.impls import sdpa as sdpa_mod  # noqa: E402
from vllm_metal_b200.attention.impls.sdpa import sdpa_forward  # noqa: E402

H, KVH, D, HIDDEN = 4, 2, 8, 32


class _Spy:
    """Records what reaches reshape_and_cache / paged_attention_primitive; `out` is filled with a ramp."""

    def __init__(self) -> None:
        self.calls: list[SimpleNamespace] = []

    def reshape_and_cache(self, key, value, key_cache, value_cache, slot_mapping):
        self.calls.append(SimpleNamespace(slot_mapping=slot_mapping.tolist(), k_shape=tuple(key.shape)))
        return key_cache, value_cache

    def paged_attention_primitive(self, query, key_cache, value_cache, num_kv_heads, scale, softcap, block_tables, seq_lens,
                                  cu_seqlens, block_size, max_seq_len, sliding_window, out, window_seqlen_q=1, sinks=None):
        c = self.calls[-1]
        c.block_tables, c.block_size, c.num_kv_heads = block_tables.tolist(), block_size, num_kv_heads
        c.q_shape, c.window = tuple(query.shape), sliding_window
        out.copy_(torch.arange(out.numel(), dtype=torch.float32).reshape(out.shape).to(out.dtype))


class _Lin(nn.Module):
    def __init__(self, i, o):
        super().__init__()
        self.w = nn.Parameter(torch.randn(o, i) * 0.1)

    def forward(self, x):
        return x @ self.w.t()


def _inner(**extra):
    m = SimpleNamespace(n_heads=H, n_kv_heads=KVH, scale=D ** -0.5, q_proj=_Lin(HIDDEN, H * D), k_proj=_Lin(HIDDEN, KVH * D),
                        v_proj=_Lin(HIDDEN, KVH * D), o_proj=lambda out: out, rope=None, q_norm=None, k_norm=None, sinks=None)
    for k, v in extra.items():
        setattr(m, k, v)
    return m


def test_mixed_batch_routes_slots_and_page_tables_by_layer_group():
    """A full-attention layer (group 0, 32-token pages) and a sliding-window layer (group 1, 16-token pages) each
    read their own scheduler group's metadata -- the reference's literals."""
    layout = MHAKVCacheLayout(num_blocks=11, tensor_sizes=(1, 1),
                              layers=(MHALayerKVLayout(0, 0, 32, KVH, D, -1), MHALayerKVLayout(1, 1, 16, KVH, D, 1024)),
                              group_block_sizes=(32, 16), slot_layers=((0,), (1,)))
    cache = B200PagedKVCache.from_layout(layout, torch.float32, "cpu")
    assert [cache.group_index_for_layer(i) for i in (0, 1)] == [0, 1]
    assert [cache.block_size_for_layer(i) for i in (0, 1)] == [32, 16]
    assert cache.key_caches[0].shape == (11, 32, KVH, D) and cache.key_caches[1].shape == (11, 16, KVH, D)
    ctx_mod.prepare_grouped([([[3], [8, 9]], 17, 1)], [([[4], [10]], 2, 0)], (32, 16))
    ctx = ctx_mod.get_context()
    spy = _Spy()
    try:
        with patch.object(sdpa_mod, "get_ops", return_value=spy), patch.object(sdpa_mod, "_FUSE_NORM_ROPE_CACHE", False):
            x = torch.ones((3, HIDDEN))
            sdpa_forward(_inner(), x, ctx, cache, 0)
            sdpa_forward(_inner(), x, ctx, cache, 1)
    finally:
        ctx_mod.clear_context()
    full, sliding = spy.calls
    assert full.slot_mapping == [3 * 32 + 17, 4 * 32, 4 * 32 + 1]
    assert full.block_tables == [[3], [4]] and full.block_size == 32 and full.window == -1
    assert sliding.slot_mapping == [9 * 16 + 1, 10 * 16, 10 * 16 + 1]
    assert sliding.block_tables == [[8, 9], [10, 0]] and sliding.block_size == 16 and sliding.window == 1024


def test_layout_from_vllm_kv_cache_config_and_runtime_adopts_it():
    from vllm.v1.kv_cache_interface import (FullAttentionSpec, KVCacheConfig, KVCacheGroupSpec, KVCacheTensor,
                                            SlidingWindowSpec)

    from vllm_metal_b200.attention.runtime.mha import MHAPagedAttentionRuntime
    from vllm_metal_b200.attention.runtime.protocol import PagedAttentionRuntime

    names = tuple(f"layers.{i}.self_attn" for i in range(4))
    full = FullAttentionSpec(block_size=32, num_kv_heads=KVH, head_size=D, dtype=torch.bfloat16)
    slide = SlidingWindowSpec(block_size=16, num_kv_heads=KVH, head_size=D, dtype=torch.bfloat16, sliding_window=128)
    nb = 6
    cfg = KVCacheConfig(
        num_blocks=nb,
        kv_cache_tensors=[KVCacheTensor(size=nb * (full if i % 2 else slide).page_size_bytes, shared_by=[names[i]])
                          for i in range(4)],
        kv_cache_groups=[KVCacheGroupSpec([names[1], names[3]], full), KVCacheGroupSpec([names[0], names[2]], slide)])
    layout = MHAKVCacheLayout.from_config(cfg, names)
    assert layout.group_block_sizes == (32, 16)
    assert [l.group_index for l in layout.layers] == [1, 0, 1, 0]
    assert [l.sliding_window for l in layout.layers] == [128, -1, 128, -1]
    assert [l.block_size for l in layout.layers] == [16, 32, 16, 32]
    assert layout.slot_layers == ((0,), (1,), (2,), (3,)) and layout.total_bytes == sum(t.size for t in cfg.kv_cache_tensors)
    with pytest.raises(ValueError, match="same layers"):
        MHAKVCacheLayout.from_config(cfg, names[:3])
    bad = KVCacheConfig(num_blocks=nb, kv_cache_tensors=[KVCacheTensor(size=7, shared_by=[names[0]])],
                        kv_cache_groups=[KVCacheGroupSpec([names[0]], full)])
    with pytest.raises(ValueError, match="does not match"):
        MHAKVCacheLayout.from_config(bad, names[:1])

    rt = MHAPagedAttentionRuntime(num_layers=4, num_kv_heads=KVH, head_dim=D, block_size=16, dtype=torch.float32, device="cpu")
    assert isinstance(rt, PagedAttentionRuntime)
    with pytest.raises(RuntimeError, match="before initialize"):
        rt.num_blocks()
    rt.adopt_layout(layout)
    assert rt.num_blocks() == nb and rt.kv_group_block_sizes() == (32, 16) and rt.kv_scheduler_group_indices() == (0, 1)
    assert rt.kv_cache.sliding_window_per_layer == [128, -1, 128, -1] and not rt.needs_step_context()


def test_patch_model_wraps_attention_and_wrapper_falls_back_without_context():
    from vllm_metal_b200.attention.impls.sdpa_wrapper import SDPAPagedAttentionWrapper, patch_sdpa_attention
    from vllm_metal_b200.attention.patching import find_attn_attr, walk_and_wrap
    from vllm_metal_b200.attention.runtime.mha import MHAPagedAttentionRuntime

    class Attn(nn.Module):
        def __init__(self):
            super().__init__()
            self.n_heads, self.n_kv_heads, self.scale = H, KVH, D ** -0.5
            self.q_proj, self.k_proj, self.v_proj = _Lin(HIDDEN, H * D), _Lin(HIDDEN, KVH * D), _Lin(HIDDEN, KVH * D)
            self.o_proj = _Lin(H * D, HIDDEN)
            self.rope = self.q_norm = self.k_norm = self.sinks = None
            self.is_local = True

        def forward(self, x, mask=None, cache=None):
            return x * 2  # "dense" path marker

    class Layer(nn.Module):
        def __init__(self):
            super().__init__()
            self.self_attn = Attn()

    class Model(nn.Module):
        def __init__(self):
            super().__init__()
            self.model = nn.Module()
            self.model.layers = nn.ModuleList([Layer(), Layer(), Layer()])

    m = Model()
    rt = MHAPagedAttentionRuntime(num_layers=3, num_kv_heads=KVH, head_dim=D, block_size=16, dtype=torch.float32, device="cpu")
    with pytest.raises(RuntimeError):
        rt.patch_model(m)
    rt.initialize(8)
    assert rt.patch_model(m) == 3
    w = m.model.layers[1].self_attn
    assert isinstance(w, SDPAPagedAttentionWrapper) and w.is_local is True and w.rope is None
    assert w._mk_layer_idx == 1 and w._mk_cache_idx == 1 and w._mk_kv_cache is rt.kv_cache
    # no paged context -> the original attention
    x = torch.ones((2, HIDDEN))
    assert torch.equal(w(x), x * 2)
    # with a context -> sdpa_forward over the spy ops, against cache layer `cache_idx`
    spy = _Spy()
    ctx_mod.prepare_unified([([5], 3)], [([6], 2, 0)], 16)
    try:
        with patch.object(sdpa_mod, "get_ops", return_value=spy), patch.object(sdpa_mod, "_FUSE_NORM_ROPE_CACHE", False):
            y = w(torch.ones((3, HIDDEN)))
    finally:
        ctx_mod.clear_context()
    assert y.shape == (3, HIDDEN) and spy.calls[0].slot_mapping == [5 * 16 + 3, 6 * 16, 6 * 16 + 1]
    # re-patching rebinds the cache in place (a re-initialised runtime), compact cache indices via cache_idx_map
    cache2 = B200PagedKVCache(2, KVH, D, 4, 16, torch.float32, "cpu")
    assert patch_sdpa_attention(m, cache2, 16, cache_idx_map={0: 0, 2: 1}, only_layers=[0, 2]) == 2
    assert m.model.layers[2].self_attn is not None and m.model.layers[2].self_attn._mk_cache_idx == 1
    assert m.model.layers[2].self_attn._mk_kv_cache is cache2 and m.model.layers[1].self_attn._mk_kv_cache is rt.kv_cache
    assert find_attn_attr(m.model.layers[0]) == "self_attn" and find_attn_attr(nn.Module()) is None
    with pytest.raises(RuntimeError, match="returned None"):
        walk_and_wrap(m, lambda i, a: None)


def test_head_dim_pad_and_output_gates():
    """Cache wider than the layer's head_dim: q/k/v are zero-padded into the kernel and the output stripped; the
    sigmoid (q_proj-split) and softplus(g_proj) gates multiply the attention output before o_proj."""
    cache = B200PagedKVCache(1, KVH, 2 * D, 4, 16, torch.float32, "cpu")  # cache head_dim 16, layer head_dim 8
    ctx_mod.prepare_unified([], [([1], 2, 0)], 16)
    ctx = ctx_mod.get_context()
    spy = _Spy()
    x = torch.randn((2, HIDDEN))
    try:
        with patch.object(sdpa_mod, "get_ops", return_value=spy), patch.object(sdpa_mod, "_FUSE_NORM_ROPE_CACHE", False):
            plain = sdpa_forward(_inner(), x, ctx, cache, 0)
            assert spy.calls[-1].q_shape == (2, H, 2 * D) and spy.calls[-1].k_shape == (2, KVH, 2 * D)
            ramp = torch.arange(2 * H * 2 * D, dtype=torch.float32).reshape(2, H, 2 * D)[..., :D].reshape(2, H * D)
            assert torch.equal(plain, ramp)  # o_proj is the identity here: the stripped kernel output
            g_proj = _Lin(HIDDEN, H)
            gated = sdpa_forward(_inner(gating=True, g_proj=g_proj), x, ctx, cache, 0)
            want = (ramp.reshape(2, H, D) * torch.nn.functional.softplus(g_proj(x)).unsqueeze(-1)).reshape(2, H * D)
            assert torch.allclose(gated, want)
            qg = _Lin(HIDDEN, 2 * H * D)
            sig = sdpa_forward(_inner(attn_output_gate=True, q_proj=qg), x, ctx, cache, 0)
            gate = qg(x).reshape(2, H, 2 * D)[..., D:].reshape(2, H * D)
            assert torch.allclose(sig, ramp * torch.sigmoid(gate))
            small = B200PagedKVCache(1, KVH, D // 2, 4, 16, torch.float32, "cpu")
            with pytest.raises(ValueError, match="exceeds the cache head_dim"):
                sdpa_forward(_inner(), x, ctx, small, 0)
    finally:
        ctx_mod.clear_context()
