"""Host-side plugin surface (B1 platform, B2 worker, quant-method hook) on CPU.

These mirror what the reference's own tests check about its plugin layer without a device
(tests/test_platform.py, tests/test_block_size_translation.py): entry points resolve, the
synthetic backend advertises MultipleOf(16), the marker quant config registers, the worker
produces the reference's layer names and per-layer spec types.  No kernels are called.
"""

from __future__ import annotations

import types

import pytest
import torch

vllm = pytest.importorskip("vllm")


def test_register_entry_points_resolve():
    import vllm_metal_b200 as pkg

    assert pkg.PLATFORM_CLS == "vllm_metal_b200.platform.B200Platform"
    # no sm_100 device in the CPU container -> vLLM must fall through to its other platforms
    assert pkg.register() in (None, pkg.PLATFORM_CLS)
    from vllm_metal_b200.platform import WORKER_CLS, B200Platform

    mod, _, cls = WORKER_CLS.rpartition(".")
    assert hasattr(__import__(mod, fromlist=[cls]), cls)
    assert B200Platform.device_type == "cuda" and B200Platform.dispatch_key == "CUDA"
    assert B200Platform.get_attn_backend_cls(None, types.SimpleNamespace(use_sparse=False, use_mla=False)) \
        == "vllm_metal_b200.attention.synthetic_backend.B200Backend"
    with pytest.raises(NotImplementedError):
        B200Platform.get_attn_backend_cls(None, types.SimpleNamespace(use_sparse=True))


def test_pyproject_declares_the_reference_entry_point_groups(monkeypatch):
    """vLLM finds the plugin through the two groups the reference declares (pyproject.toml:104-112); each
    target must resolve to a callable, and register() stays inert without the operator's opt-in."""
    import importlib
    import os
    import tomllib

    import vllm_metal_b200 as pkg

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    with open(os.path.join(root, "pyproject.toml"), "rb") as f:
        eps = tomllib.load(f)["project"]["entry-points"]
    assert set(eps) == {"vllm.platform_plugins", "vllm.general_plugins"}
    for group in eps.values():
        for target in group.values():
            mod, _, fn = target.partition(":")
            assert callable(getattr(importlib.import_module(mod), fn))
    monkeypatch.delenv("VLLM_B200_PLUGIN", raising=False)
    assert pkg.register() is None
    monkeypatch.setenv("VLLM_B200_PLUGIN", "1")
    monkeypatch.setattr(pkg, "_cuda_sm100_available", lambda: True)
    assert pkg.register() == pkg.PLATFORM_CLS


def test_synthetic_backend_advertises_multiple_of_16_and_refuses_dispatch():
    from vllm.v1.attention.backend import MultipleOf

    from vllm_metal_b200.attention.synthetic_backend import B200Backend

    sizes = B200Backend.get_supported_kernel_block_sizes()
    assert len(sizes) == 1 and isinstance(sizes[0], MultipleOf) and sizes[0].base == 16
    for fn in (B200Backend.get_impl_cls, B200Backend.get_builder_cls, B200Backend.get_kv_cache_shape):
        with pytest.raises(NotImplementedError):
            fn()


def test_platform_config_gates():
    from vllm_metal_b200.platform import WORKER_CLS, B200Platform

    def cfg(tp=1, pp=1, block=None, user=False):
        return types.SimpleNamespace(
            parallel_config=types.SimpleNamespace(worker_cls="auto", tensor_parallel_size=tp,
                                                  pipeline_parallel_size=pp, disable_custom_all_reduce=False),
            cache_config=types.SimpleNamespace(block_size=block, user_specified_block_size=user),
            compilation_config=None)

    c = cfg(tp=2)
    B200Platform.check_and_update_config(c)
    assert c.parallel_config.worker_cls == WORKER_CLS and c.parallel_config.disable_custom_all_reduce
    assert c.cache_config.block_size == 16
    with pytest.raises(ValueError):
        B200Platform.check_and_update_config(cfg(pp=2))
    with pytest.raises(ValueError):
        B200Platform.check_and_update_config(cfg(tp=3))
    c = cfg(block=544, user=True)  # the reference's hybrid example: translated 544 -> 32 x 17
    B200Platform.update_block_size_for_backend(c)
    assert c.cache_config.block_size == 544
    with pytest.raises(ValueError):
        B200Platform.update_block_size_for_backend(cfg(block=7, user=True))


def test_validate_request_rejects_logits_processor_controls():
    from vllm.sampling_params import SamplingParams

    from vllm_metal_b200.platform import B200Platform

    B200Platform.validate_request(None, SamplingParams(temperature=0.0))
    for kw in (dict(min_p=0.1), dict(logit_bias={1: 2.0}), dict(min_tokens=3, max_tokens=8)):
        with pytest.raises(Exception, match="logits"):
            B200Platform.validate_request(None, SamplingParams(**kw))


def test_marker_quant_config_registers_idempotently():
    import vllm_metal_b200 as pkg
    from vllm.model_executor.layers.quantization import get_quantization_config

    pkg.register_quant_config()
    pkg.register_quant_config()
    cls = get_quantization_config("mlx_affine")
    c = cls.from_config({"quantization": {"bits": 8, "group_size": 32}})
    assert (c.bits, c.group_size) == (8, 32) and cls.get_name() == "mlx_affine"
    with pytest.raises(ValueError):
        cls(bits=3)
    with pytest.raises(ValueError):
        cls(group_size=48)


def test_worker_kv_cache_spec_names_and_window_types():
    from vllm.v1.kv_cache_interface import FullAttentionSpec, SlidingWindowSpec

    from vllm_metal_b200.v1.worker import B200Worker, model_args_from_hf_config

    hf = types.SimpleNamespace(num_attention_heads=8, hidden_size=512, num_key_value_heads=4, head_dim=64,
                               intermediate_size=1024, vocab_size=1000, num_hidden_layers=4, rms_norm_eps=1e-5,
                               rope_theta=150000.0, tie_word_embeddings=False, model_type="gpt_oss",
                               sliding_window=128,
                               layer_types=["sliding_attention", "full_attention"] * 2,
                               quantization={"bits": 4, "group_size": 32})
    args = model_args_from_hf_config(hf)
    assert (args.num_layers, args.n_kv_heads, args.head_dim, args.group_size) == (4, 4, 64, 32)
    assert args.sliding_window_per_layer() == [128, -1, 128, -1]
    w = object.__new__(B200Worker)
    w.model = types.SimpleNamespace(args=args)
    w.parallel_config = types.SimpleNamespace(tensor_parallel_size=2)
    w.cache_config = types.SimpleNamespace(block_size=16)
    specs = w.get_kv_cache_spec()
    assert list(specs) == [f"layers.{i}.self_attn" for i in range(4)]  # cache_policy.py:383,406-411
    assert isinstance(specs["layers.0.self_attn"], SlidingWindowSpec)
    assert isinstance(specs["layers.1.self_attn"], FullAttentionSpec)
    assert specs["layers.1.self_attn"].num_kv_heads == 2  # sharded over TP=2
    assert w.get_cache_block_size_bytes() == 2 * 4 * 16 * 2 * 64 * 2


def test_worker_packs_decodes_before_prefills():
    from vllm_metal_b200.v1.worker import B200Worker

    w = object.__new__(B200Worker)
    so = types.SimpleNamespace(decodes=[(11, [3, 4], 20), (12, [5], 7)], prefills=[([1, 2, 3], [6], 0)])
    tokens, dec, pre = w._pack(so)
    assert tokens == [11, 12, 1, 2, 3]
    assert dec == [([3, 4], 20), ([5], 7)] and pre == [([6], 3, 0)]
