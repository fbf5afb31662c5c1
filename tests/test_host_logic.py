"""Host-side product logic on CPU: per-step context, block-size translation,
quantiser, module protocol, and the C-ABI library's exported symbols.

These read like the reference's own tests (tests/test_paged_attention.py TestPrepare,
tests/test_block_size_translation.py) but run against ``vllm_metal_b200``.
"""

from __future__ import annotations

import ctypes
import os
import re

import numpy as np
import pytest
import torch

from oracle import affine_quant as oq
from oracle import metadata as ometa
from vllm_metal_b200.attention.context import (
    clear_context,
    get_context,
    prepare_grouped,
    prepare_unified,
)
from vllm_metal_b200.attention.impls import sdpa as sdpa_mod
from vllm_metal_b200.metal import EXPORTED_SYMBOLS, PagedAttnArgs, library_path
from vllm_metal_b200.quant import affine as paffine

from .test_oracle_golden import LITERAL_PREPARE

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(autouse=True)
def _clear():
    yield
    clear_context()


@pytest.mark.parametrize("case", LITERAL_PREPARE)
def test_prepare_unified_literals(case):
    dec, pre, bs, merge, want = case
    prepare_unified(dec, pre, bs, merge_verify_windows=merge)
    ctx = get_context()
    assert ctx is not None
    for k, v in want.items():
        assert getattr(ctx, k) == v, k


def test_prepare_grouped_builds_distinct_group_metadata():
    prepare_grouped([([[3, 4], [8]], 17, 1)], [([[5, 6], [9]], 2, 0)], (16, 32))
    ctx = get_context()
    assert ctx.kv_groups[0].slot_mapping == [4 * 16 + 1, 5 * 16, 5 * 16 + 1]
    assert ctx.kv_groups[1].slot_mapping == [8 * 32 + 17, 9 * 32, 9 * 32 + 1]
    assert ctx.kv_groups[0].block_tables == [[3, 4], [5, 6]]
    assert ctx.kv_groups[1].block_tables == [[8], [9]]
    assert ctx.slot_mapping == ctx.kv_groups[0].slot_mapping
    assert ctx.block_tables == ctx.kv_groups[0].block_tables


def test_prepare_grouped_repeats_decode_table_for_each_verification_token():
    prepare_grouped([([[3], [8, 9]], 17, 2)], [], (32, 16))
    ctx = get_context()
    assert ctx.context_lens == [18, 19]
    assert ctx.kv_groups[0].block_tables == [[3], [3]]
    assert ctx.kv_groups[1].block_tables == [[8, 9], [8, 9]]


def test_prepare_matches_reference_fixtures_and_oracle(golden_metadata):
    for case in golden_metadata["prepare"]:
        dec = [tuple(r) for r in case["decode"]]
        pre = [tuple(r) for r in case["prefill"]]
        prepare_grouped(dec, pre, tuple(case["block_sizes"]), merge_verify_windows=case["merge"])
        ctx = get_context()
        want = case["out"]
        for k in ("slot_mapping", "block_tables", "context_lens", "cu_seqlens", "offsets",
                  "num_decode_requests", "verify_window_q"):
            assert getattr(ctx, k) == want[k], k
        for g, wg in zip(ctx.kv_groups, want["kv_groups"], strict=True):
            assert g.slot_mapping == wg["slot_mapping"]
            assert g.block_tables == wg["block_tables"]
            assert g.block_size == wg["block_size"]
        clear_context()


def test_pick_kernel_block_size():
    for bs in (32, 16, 8):
        assert sdpa_mod._pick_kernel_block_size(bs) == bs
    assert sdpa_mod._pick_kernel_block_size(544) == 32
    assert sdpa_mod._pick_kernel_block_size(48) == 16
    assert sdpa_mod._pick_kernel_block_size(24) == 8
    with pytest.raises(ValueError, match="not divisible"):
        sdpa_mod._pick_kernel_block_size(7)


def test_build_block_tables_matches_fixtures(golden_metadata):
    for case in golden_metadata["build_block_tables"]:
        bt, kbs = sdpa_mod._build_block_tables_np(case["raw"], case["cache_block_size"])
        assert kbs == case["kernel_block_size"]
        assert bt.dtype == np.int32 and bt.tolist() == case["tables"]
    bt, kbs = sdpa_mod._build_block_tables_np([], 544)
    assert bt.shape == (0, 0) and kbs == 544
    bt, kbs = sdpa_mod._build_block_tables_np([[0, 1], [2]], 544)
    assert bt[1].tolist() == list(range(34, 51)) + list(range(17))


def test_positions_follow_offsets():
    prepare_unified([([5, 6], 7)], [([10, 11], 5, 0), ([1, 2], 3, 4)], block_size=4)
    ctx = get_context()
    assert sdpa_mod._positions(ctx).tolist() == [7, 0, 1, 2, 3, 4, 4, 5, 6]


# ------------------------------------------------------------- quantiser ----

@pytest.mark.parametrize("bits,group", [(4, 32), (4, 64), (4, 128), (8, 64)])
def test_torch_quantize_bit_exact_vs_oracle(bits, group):
    g = torch.Generator().manual_seed(0)
    w = (torch.randn(24, 256, generator=g) * 0.02).to(torch.bfloat16)
    wq, s, b = paffine.quantize(w, group, bits, torch.bfloat16)
    owq, os_, ob = oq.quantize(w.float().numpy(), group, bits, "bfloat16")
    assert np.array_equal(wq.numpy().view(np.uint32), owq)
    assert np.array_equal(s.float().numpy(), os_)
    assert np.array_equal(b.float().numpy(), ob)
    assert np.array_equal(paffine.unpack(wq, bits).numpy(), oq.unpack(owq, bits))
    deq = paffine.dequantize_reference(wq, s, b, group, bits).numpy()
    assert np.array_equal(deq, oq.dequantize(owq, os_, ob, group, bits))


def test_quantized_linear_module_protocol():
    """Attribute names/shapes the reference duck-types (quant/awq_loader.py:165-180,
    v1/lora/layers.py:21-50)."""
    from vllm_metal_b200.quant.linear import QuantizedLinear

    w = torch.randn(32, 128) * 0.02
    m = QuantizedLinear.from_dense(w.to(torch.bfloat16), None, 64, 4)
    assert hasattr(m, "bits") and hasattr(m, "group_size")
    assert m.weight.shape == (32, 128 * 4 // 32) and m.weight.dtype == torch.int32
    assert m.scales.shape == (32, 2) == m.biases.shape
    assert m.in_features == m.scales.shape[1] * m.group_size == 128
    assert m.out_features == m.scales.shape[0] == 32
    a = m.shard_rows(1, 2)
    assert a.weight.shape == (16, 16) and torch.equal(a.weight, m.weight[16:])
    c = m.shard_cols(1, 2)
    assert c.weight.shape == (32, 8) and torch.equal(c.scales, m.scales[:, 1:])
    with pytest.raises(ValueError):
        m.shard_cols(0, 4)  # 128/4 = 32 is not a multiple of group 64
    with pytest.raises(RuntimeError, match="CUDA"):
        m(torch.zeros(1, 128, dtype=torch.bfloat16))  # no CPU fallback


# ------------------------------------------------------------------ C-ABI ----

def _declared_symbols() -> set[str]:
    names = set()
    for h in ("b200_paged_ops.h", "b200_decoder_ops.h", "b200_tp_ops.h"):
        path = os.path.join(ROOT, "include", h)
        if not os.path.exists(path):
            continue
        text = re.sub(r"/\*.*?\*/", "", open(path).read(), flags=re.S)
        names |= set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", text))
    return names


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(library_path())
    declared = _declared_symbols()
    assert declared, "no declarations parsed"
    assert declared == set(EXPORTED_SYMBOLS)
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in include/*.h but not exported"
    assert lib.b200_abi_version() == 3  # round 2: x_dtype / out_dtype arguments, args.out_dtype, args.window_seqlen_q
    assert lib.b200_partition_size() == 512


def test_args_struct_matches_header_layout():
    # 10 pointers/size_t + 12 int32 + 6 int64 + 2 float + 2 int32 (out_dtype, window_seqlen_q)
    assert ctypes.sizeof(PagedAttnArgs) == 10 * 8 + 12 * 4 + 6 * 8 + 2 * 4 + 4 + 4
    assert PagedAttnArgs.out_dtype.offset == 10 * 8 + 12 * 4 + 6 * 8 + 2 * 4
    assert PagedAttnArgs.window_seqlen_q.offset == PagedAttnArgs.out_dtype.offset + 4
    assert PagedAttnArgs.q_stride.offset == 10 * 8 + 12 * 4
    assert PagedAttnArgs.scale.offset == 10 * 8 + 12 * 4 + 6 * 8


def test_ops_refuse_cpu_tensors():
    """No CPU fallback: the reference's eval_cpu throws (paged_ops.cpp:740-743,1121-1123)."""
    from vllm_metal_b200.metal import get_ops

    ops = get_ops()  # loading needs no GPU
    k = torch.zeros(2, 2, 64, dtype=torch.bfloat16)
    kc = torch.zeros(2, 16, 2, 64, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError, match="CUDA"):
        ops.reshape_and_cache(k, k, kc, kc, torch.zeros(2, dtype=torch.int64))
    with pytest.raises(RuntimeError, match="CUDA"):
        ops.paged_attention_primitive(k, kc, kc, 2, 1.0, 0.0, torch.zeros(1, 1, dtype=torch.int32),
                                      torch.zeros(1, dtype=torch.int32), torch.zeros(2, dtype=torch.int32),
                                      16, 16, -1, k)


def test_oracle_not_imported_by_product():
    """The product package must never import the oracle."""
    pkg = os.path.join(ROOT, "vllm_metal_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f


def test_oracle_moe_routing_and_activation_semantics():
    """oracle/moe.py: gpt-oss routing = softmax over the k selected logits; Qwen3-MoE routing = softmax over all, top-k,
    renormalised; ties go to the lowest expert id; the clipped SwiGLU saturates at the limit."""
    from oracle import moe

    logits = np.array([[1.0, 3.0, 3.0, -2.0], [0.0, 0.0, 0.0, 0.0]], np.float32)
    idx, sc = moe.route(logits, 2, "topk_softmax")
    assert idx.tolist() == [[1, 2], [0, 1]]
    np.testing.assert_allclose(sc, [[0.5, 0.5], [0.5, 0.5]], atol=1e-6)
    idx2, sc2 = moe.route(logits, 2, "softmax_topk", norm_topk_prob=False)
    assert idx2.tolist() == [[1, 2], [0, 1]]
    p = np.exp(logits[0] - 3.0) / np.exp(logits[0] - 3.0).sum()
    np.testing.assert_allclose(sc2[0], p[[1, 2]], rtol=1e-6)
    np.testing.assert_allclose(moe.route(logits, 2, "softmax_topk")[1].sum(-1), 1.0, atol=1e-6)
    g, u = np.array([100.0, -100.0, 1.0], np.float32), np.array([100.0, -100.0, 0.5], np.float32)
    a = moe.activation(g, u, "gpt_oss")
    np.testing.assert_allclose(a[0], 7.0 / (1.0 + np.exp(-1.702 * 7.0)) * 8.0, rtol=1e-6)  # both clipped at the limit
    assert abs(a[1]) < 1e-30                                                            # sigmoid(-170) kills it
    np.testing.assert_allclose(a[2], 1.0 / (1.0 + np.exp(-1.702)) * 1.5, rtol=1e-6)
    np.testing.assert_allclose(moe.activation(g[2:], u[2:], "silu"), 1.0 / (1.0 + np.exp(-1.0)) * 0.5, rtol=1e-6)


def test_switch_linear_shards_are_slices_of_the_stacked_experts():
    """Tensor-parallel MoE: gate / up shard every expert's output rows, down shards K on group boundaries with the
    dense bias on rank 0 only -- literal slices of the packed tensors, and the rank partials add up to the whole."""
    import torch

    from oracle import affine_quant as oq
    from vllm_metal_b200.quant.switch_linear import QuantizedSwitchLinear

    g = torch.Generator().manual_seed(3)
    sw = QuantizedSwitchLinear.from_dense(torch.randn((3, 64, 256), generator=g) * 0.1, torch.randn((3, 64), generator=g),
                                          64, 4, torch.bfloat16)
    up = [sw.shard_rows(r, 2) for r in range(2)]
    assert torch.equal(torch.cat([u.weight for u in up], 1), sw.weight)
    assert torch.equal(torch.cat([u.bias for u in up], 1), sw.bias)
    down = [sw.shard_cols(r, 2) for r in range(2)]
    assert down[0].bias is not None and down[1].bias is None and down[0].in_features == 128
    assert torch.equal(torch.cat([d.scales for d in down], 2), sw.scales)

    def dense(m, e):
        return oq.dequantize(m.weight[e].numpy().view(np.uint32), m.scales[e].float().numpy(), m.biases[e].float().numpy(),
                             m.group_size, m.bits).astype(np.float64)

    x = torch.randn((5, 256), generator=g).double().numpy()
    for e in range(3):
        full = x @ dense(sw, e).T
        part = sum(x[:, r * 128:(r + 1) * 128] @ dense(down[r], e).T for r in range(2))
        np.testing.assert_allclose(part, full, rtol=1e-10, atol=1e-10)
    with pytest.raises(ValueError):
        sw.shard_cols(0, 8)  # 256 / 8 = 32 is not a multiple of the group size
