"""Quantized MoE expert path through the C-ABI linears vs oracle/moe.py."""

from __future__ import annotations

import numpy as np
import pytest
import torch

from oracle import moe as omoe
from tests.gpu_helpers import to_np

pytestmark = pytest.mark.gpu


def _stack(m):
    return (m.weight.cpu().numpy().view(np.uint32), to_np(m.scales), to_np(m.biases),
            None if m.bias is None else to_np(m.bias))


def _build(hidden, inter, experts, top_k, bits, group, bias, act, routing, seed):
    from vllm_metal_b200.quant.switch_linear import MoEBlock, QuantizedSwitchLinear, SwitchGLU

    g = torch.Generator(device="cuda").manual_seed(seed)
    rnd = lambda *s: torch.randn(s, device="cuda", generator=g)  # noqa: E731
    glu = SwitchGLU(hidden, inter, experts, bias, group, bits, torch.bfloat16, "cuda", activation=act)
    for name, (n, k) in (("gate_proj", (inter, hidden)), ("up_proj", (inter, hidden)), ("down_proj", (hidden, inter))):
        dense = rnd(experts, n, k) * (0.7 / k ** 0.5)  # activations stay O(1): the bound is the linears' north-star tolerance
        b = rnd(experts, n) * 0.1 if bias else None
        setattr(glu, name, QuantizedSwitchLinear.from_dense(dense, b, group, bits, torch.bfloat16))
    router = torch.nn.Linear(hidden, experts, bias=bias, device="cuda", dtype=torch.bfloat16).requires_grad_(False)
    with torch.no_grad():
        router.weight.copy_(rnd(experts, hidden) * 0.5)
        if bias:
            router.bias.copy_(rnd(experts) * 0.1)
    return MoEBlock(router, glu, top_k, routing)


@pytest.mark.parametrize("bits,group", [(4, 64), (8, 32)])
@pytest.mark.parametrize("act,routing,bias", [("gpt_oss", "topk_softmax", True), ("silu", "softmax_topk", False)])
@pytest.mark.parametrize("tokens", [1, 37])
def test_moe_block_vs_oracle(bits, group, act, routing, bias, tokens):
    hidden, inter, experts, top_k = 256, 384, 8, 2
    blk = _build(hidden, inter, experts, top_k, bits, group, bias, act, routing, seed=tokens + bits)
    x = torch.randn((tokens, hidden), device="cuda", generator=torch.Generator(device="cuda").manual_seed(9)).bfloat16()
    y = blk(x)
    idx, scores = blk.route(x)
    torch.cuda.synchronize()
    glu = blk.experts
    rw, rb = to_np(blk.router.weight), None if blk.router.bias is None else to_np(blk.router.bias)
    # routing parity: the oracle's own top-k picks the same experts (as sets; random logits do not tie)
    _, oidx, oscores = omoe.moe_block(to_np(x), rw, rb, _stack(glu.gate_proj), _stack(glu.up_proj), _stack(glu.down_proj),
                                      top_k, group, bits, "bfloat16", routing, act)
    same = [set(a) == set(b) for a, b in zip(idx.cpu().tolist(), oidx.tolist(), strict=True)]
    assert np.mean(same) >= 0.95  # a bf16-rounded logit pair can swap at the k-th place
    # expert path parity on the product's routing: north-star bound of the quantized linears
    want, _, _ = omoe.moe_block(to_np(x), rw, rb, _stack(glu.gate_proj), _stack(glu.up_proj), _stack(glu.down_proj),
                                top_k, group, bits, "bfloat16", routing, act,
                                indices=idx.cpu().numpy(), scores=scores.cpu().numpy())
    got = to_np(y)
    assert np.isfinite(got).all()
    err = np.abs(got - want)
    assert (err <= 2e-2 + 1e-2 * np.abs(want)).all(), f"max err {err.max()} at |want| max {np.abs(want).max()}"
    np.testing.assert_allclose(scores.cpu().numpy().sum(-1), 1.0, atol=1e-5)


def test_switch_linear_gather_form_matches_per_expert_linears():
    """``QuantizedSwitchLinear(x, indices)`` (mx.gather_qmm form, x shared or per pair) = the dense quantized linear of
    each picked expert, bit for bit (same kernel, same rows)."""
    from vllm_metal_b200.metal import get_ops
    from vllm_metal_b200.quant.switch_linear import QuantizedSwitchLinear

    g = torch.Generator(device="cuda").manual_seed(3)
    sw = QuantizedSwitchLinear.from_dense(torch.randn((6, 192, 128), device="cuda", generator=g) * 0.1,
                                          torch.randn((6, 192), device="cuda", generator=g) * 0.1, 64, 4)
    x = torch.randn((9, 128), device="cuda", generator=g).bfloat16()
    idx = torch.randint(0, 6, (9, 3), device="cuda", generator=g)
    y = sw(x, idx)
    assert y.shape == (9, 3, 192)
    ops = get_ops()
    for t in range(9):
        for j in range(3):
            e = int(idx[t, j])
            ref = ops.quantized_matmul(x[t:t + 1], sw.weight[e], sw.scales[e], sw.biases[e], group_size=64, bits=4,
                                       bias=sw.bias[e])
            assert torch.equal(y[t, j], ref[0]), (t, j, e)
    x3 = torch.randn((9, 3, 128), device="cuda", generator=g).bfloat16()
    y3 = sw(x3, idx)
    e = int(idx[4, 2])
    ref = ops.quantized_matmul(x3[4, 2:3], sw.weight[e], sw.scales[e], sw.biases[e], group_size=64, bits=4, bias=sw.bias[e])
    assert torch.equal(y3[4, 2], ref[0])


def test_moe_at_the_gpt_oss_expert_shape():
    """gpt-oss-20b's expert block: hidden 2880, 32 experts, top-4, int4 g64 (K = 2880 = 45 x 64), bs 32 decode rows."""
    hidden, inter, experts, top_k = 2880, 2880, 32, 4
    blk = _build(hidden, inter, experts, top_k, 4, 64, True, "gpt_oss", "topk_softmax", seed=5)
    x = torch.randn((32, hidden), device="cuda", generator=torch.Generator(device="cuda").manual_seed(1)).bfloat16()
    y = blk(x)
    idx, scores = blk.route(x)
    glu = blk.experts
    want, _, _ = omoe.moe_block(to_np(x), to_np(blk.router.weight), to_np(blk.router.bias), _stack(glu.gate_proj),
                                _stack(glu.up_proj), _stack(glu.down_proj), top_k, 64, 4, "bfloat16", "topk_softmax",
                                "gpt_oss", indices=idx.cpu().numpy(), scores=scores.cpu().numpy())
    err = np.abs(to_np(y) - want)
    assert (err <= 2e-2 + 1e-2 * np.abs(want)).all(), f"max err {err.max()}"
