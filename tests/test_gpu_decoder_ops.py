"""GPU parity of the ops either side of the hot path (include/b200_decoder_ops.h).

* RMSNorm (vectorised and scalar routes, fused residual add) vs the oracle.
* The fused per-head norm -> RoPE -> cache-write kernel is BIT-IDENTICAL to the step-by-step
  sequence the reference runs (sdpa.py:348-368 norm + RoPE, :638 reshape_and_cache), and the
  sequence itself matches the oracle (kv write bit-exact, rotation within bf16 tolerance).
"""

from __future__ import annotations

import numpy as np
import pytest
import torch

from oracle import decoder as odec
from oracle import kv_write as okv
from oracle._numeric import round_to

from .gpu_helpers import TORCH_DT, assert_close, to_np

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dim", [64, 128, 520, 1024, 4096, 5120, 8192, 8200, 12288])
@pytest.mark.parametrize("dtype", ["bfloat16", "float16"])
@pytest.mark.parametrize("with_residual", [False, True])
def test_rms_norm_matches_oracle(ops, dim, dtype, with_residual):
    g = torch.Generator().manual_seed(dim)
    tdt = TORCH_DT[dtype]
    rows = 5
    x = torch.randn((rows, dim), generator=g).to(tdt).cuda()
    w = (1.0 + 0.1 * torch.randn(dim, generator=g)).to(tdt).cuda()
    res = torch.randn((rows, dim), generator=g).to(tdt).cuda() if with_residual else None
    xs = to_np(x)
    if with_residual:
        summed = round_to(xs + to_np(res), dtype)
        want = odec.rms_norm(summed, to_np(w), 1e-5, dtype)
    else:
        want = odec.rms_norm(xs, to_np(w), 1e-5, dtype)
    got = ops.rms_norm(x, w, 1e-5, residual=res)
    torch.cuda.synchronize()
    assert_close(got, want, dtype, f"rms_norm dim={dim}")
    if with_residual:  # the residual stream is updated in place, rounded once
        np.testing.assert_array_equal(to_np(res), summed)


def test_rms_norm_strided_rows(ops):
    g = torch.Generator().manual_seed(1)
    buf = torch.randn((7, 3 * 4096), generator=g).bfloat16().cuda()
    x = buf[:, 4096:8192]  # row stride 12288, 16-byte aligned: vector route
    w = torch.ones(4096, dtype=torch.bfloat16, device="cuda")
    got = ops.rms_norm(x, w, 1e-6)
    torch.cuda.synchronize()
    assert_close(got, odec.rms_norm(to_np(x), to_np(w), 1e-6, "bfloat16"), "bfloat16", "strided rms_norm")


CASES = [
    # (T, Hq, Hk, D, rot_dim, traditional, qk_norm, dtype)
    (9, 8, 2, 64, 64, False, True, "bfloat16"),     # tiny preset (Qwen3 style)
    (64, 32, 8, 128, 128, False, False, "bfloat16"),  # Llama-3.1-8B decode step
    (5, 16, 8, 128, 128, False, True, "bfloat16"),  # Qwen3-0.6B
    (7, 64, 8, 64, 64, False, False, "bfloat16"),   # gpt-oss
    (6, 4, 4, 96, 32, True, True, "float16"),       # partial rotary, interleaved pairs
    (3, 2, 1, 80, 0, False, True, "bfloat16"),      # no RoPE
    (4, 4, 2, 128, 128, False, False, "float32"),
]


@pytest.mark.parametrize("case", CASES, ids=[str(c) for c in CASES])
def test_fused_norm_rope_cache_is_bit_identical_to_the_sequence(ops, case):
    t, hq, hk, d, rot, trad, qk_norm, dtype = case
    tdt = TORCH_DT[dtype]
    g = torch.Generator().manual_seed(t * 131 + d)
    bs, nb = 16, 11
    qkv = torch.randn((t, (hq + 2 * hk) * d), generator=g).to(tdt).cuda()
    qn = (1.0 + 0.1 * torch.randn(d, generator=g)).to(tdt).cuda() if qk_norm else None
    kn = (1.0 + 0.1 * torch.randn(d, generator=g)).to(tdt).cuda() if qk_norm else None
    pos = torch.randint(0, 5000, (t,), generator=g).int().cuda()
    slots = torch.randperm(nb * bs, generator=g)[:t].long()
    slots[t // 2] = -1  # padding row: q/k still updated, nothing cached
    slots = slots.cuda()
    kc0 = torch.randn((nb, bs, hk, d), generator=g).to(tdt).cuda()
    vc0 = torch.randn((nb, bs, hk, d), generator=g).to(tdt).cuda()
    eps, base = 1e-6, 1e6

    def views(buf):
        q, k, v = buf.split([hq * d, hk * d, hk * d], dim=-1)
        return q.reshape(t, hq, d), k.reshape(t, hk, d), v.reshape(t, hk, d)

    # --- step by step (the reference's order) ---
    q1, k1, v1 = (x.contiguous() for x in views(qkv.clone()))
    if qk_norm:
        q1, k1 = ops.rms_norm(q1, qn, eps), ops.rms_norm(k1, kn, eps)
    if rot:
        ops.rope(q1, k1, pos, rot, base, trad)
    kc1, vc1 = kc0.clone(), vc0.clone()
    ops.reshape_and_cache(k1, v1, kc1, vc1, slots)

    # --- fused, on strided views of the packed qkv buffer ---
    buf = qkv.clone()
    q2, k2, v2 = views(buf)
    kc2, vc2 = kc0.clone(), vc0.clone()
    ops.norm_rope_cache(q2, k2, v2, kc2, vc2, slots, positions=pos, rot_dim=rot, base=base, traditional=trad,
                        q_norm_weight=qn, k_norm_weight=kn, eps=eps)
    torch.cuda.synchronize()
    for name, a, b in (("q", q1, q2), ("k", k1, k2), ("key_cache", kc1, kc2), ("value_cache", vc1, vc2)):
        assert torch.equal(a, b.contiguous()), f"{name} differs between fused and step-by-step"
    assert torch.equal(v2.contiguous(), views(qkv)[2].contiguous()), "v must not be modified"

    # --- and the sequence matches the oracle ---
    qs, ks, vs = (to_np(x) for x in views(qkv))
    if qk_norm:
        qs, ks = odec.rms_norm(qs, to_np(qn), eps, dtype), odec.rms_norm(ks, to_np(kn), eps, dtype)
    if rot:
        p = pos.cpu().numpy()
        qs, ks = odec.rope(qs, p, rot, base, trad, dtype), odec.rope(ks, p, rot, base, trad, dtype)
    assert_close(q2, qs, dtype, "fused q vs oracle")
    assert_close(k2, ks, dtype, "fused k vs oracle")
    # cache write: bit-exact scatter of the GPU's own k rows and of v
    kco, vco = to_np(kc0).copy(), to_np(vc0).copy()
    okv.reshape_and_cache(to_np(k2), vs, kco, vco, slots.cpu().numpy())
    np.testing.assert_array_equal(to_np(kc2), kco)
    np.testing.assert_array_equal(to_np(vc2), vco)


def test_fused_rejects_bad_arguments(ops):
    q = torch.zeros((2, 4, 64), dtype=torch.bfloat16, device="cuda")
    k = torch.zeros((2, 2, 64), dtype=torch.bfloat16, device="cuda")
    kc = torch.zeros((3, 16, 2, 64), dtype=torch.bfloat16, device="cuda")
    slots = torch.zeros(2, dtype=torch.int64, device="cuda")
    with pytest.raises(ValueError):
        ops.norm_rope_cache(q, k, k, kc, kc, slots.int())  # slot dtype
    with pytest.raises(ValueError):
        ops.norm_rope_cache(q, k, k, kc, kc, slots, rot_dim=64)  # RoPE without positions
    with pytest.raises(ValueError):
        ops.norm_rope_cache(q, k, k, kc.float(), kc.float(), slots)  # dtype mismatch
    with pytest.raises(ValueError):
        ops.norm_rope_cache(q, k, k, kc, kc, slots, positions=slots.int(), rot_dim=66)  # rot_dim > D


def test_sdpa_forward_fused_and_unfused_agree(ops):
    """The product path (fused) and the reference's step order produce identical logits."""
    import dataclasses

    from vllm_metal_b200.attention.caches.kv_cache import B200PagedKVCache
    from vllm_metal_b200.attention.impls import sdpa
    from vllm_metal_b200.lm.decoder import PRESETS, Model
    from vllm_metal_b200.v1.model_runner import PagedModelRunner

    args = dataclasses.replace(PRESETS["tiny"])
    model = Model(args, torch.bfloat16, "cuda", seed=3)
    outs = []
    for fuse in (True, False):
        sdpa._FUSE_NORM_ROPE_CACHE = fuse
        try:
            cache = B200PagedKVCache(args.num_layers, args.n_kv_heads, args.head_dim, 9, 16, torch.bfloat16, "cuda")
            runner = PagedModelRunner(model, cache, max_batch=2, max_model_len=64, use_cuda_graph=False)
            toks = runner.execute(list(range(3, 40)), [], [([1, 2, 3], 20, 0), ([5, 6], 17, 0)])
            toks += runner.execute(toks, [([1, 2, 3], 20), ([5, 6], 17)], [])
            outs.append((toks, cache.key_caches[0].clone(), cache.value_caches[1].clone()))
        finally:
            sdpa._FUSE_NORM_ROPE_CACHE = True
    assert outs[0][0] == outs[1][0]
    assert torch.equal(outs[0][1], outs[1][1]) and torch.equal(outs[0][2], outs[1][2])


@pytest.mark.parametrize("dtype", ["bfloat16", "float16", "float32"])
@pytest.mark.parametrize("rows,vocab", [(1, 7), (3, 1000), (64, 128256), (5, 151937), (2, 8), (4, 65)])
def test_argmax_is_exact_with_lowest_index_ties(ops, dtype, rows, vocab):
    """Greedy top-1 (v1/sampling_batch.py:342-348): bit-exact index, lowest index wins ties."""
    g = torch.Generator(device="cuda").manual_seed(rows * 131 + vocab)
    # few distinct values -> many exact ties, also across the cluster's slices
    x = torch.randint(-3, 4, (rows, vocab), device="cuda", generator=g).to(TORCH_DT[dtype])
    got = ops.argmax(x).cpu().numpy()
    want = np.argmax(to_np(x).astype(np.float64), axis=-1)
    np.testing.assert_array_equal(got, want)
    x = torch.randn((rows, vocab), device="cuda", generator=g).to(TORCH_DT[dtype])
    np.testing.assert_array_equal(ops.argmax(x).cpu().numpy(), np.argmax(to_np(x).astype(np.float64), axis=-1))


def test_argmax_strided_and_unaligned_rows(ops):
    g = torch.Generator(device="cuda").manual_seed(9)
    big = torch.randn((6, 4099), device="cuda", generator=g).bfloat16()
    view = big[:, 3:4000]  # row starts are not 16-byte aligned, stride != width
    np.testing.assert_array_equal(ops.argmax(view).cpu().numpy(),
                                  np.argmax(to_np(view).astype(np.float64), axis=-1))


def test_programmatic_dependent_launch_does_not_change_results(ops):
    """PDL only overlaps prologues with the previous kernel's tail: a chain of dependent kernels
    must give bit-identical results with it on and off."""
    g = torch.Generator(device="cuda").manual_seed(4)
    n, k, m = 512, 1024, 64
    w = torch.randint(-2**31, 2**31 - 1, (n, k // 8), dtype=torch.int32, device="cuda", generator=g)
    s = (torch.rand((n, k // 64), device="cuda", generator=g) * 0.01 + 0.001).bfloat16()
    b = (-8.0 * s.float()).bfloat16()
    w2 = torch.randint(-2**31, 2**31 - 1, (k, n // 8), dtype=torch.int32, device="cuda", generator=g)
    s2 = (torch.rand((k, n // 64), device="cuda", generator=g) * 0.01 + 0.001).bfloat16()
    b2 = (-8.0 * s2.float()).bfloat16()
    nw = torch.ones((k,), device="cuda", dtype=torch.bfloat16)
    x0 = torch.randn((m, k), device="cuda", generator=g).bfloat16()

    def chain():
        x = x0.clone()
        for _ in range(6):
            h = ops.rms_norm(x, nw, 1e-6)
            y = ops.quantized_matmul(h, w, s, b, group_size=64, bits=4)
            x = ops.quantized_matmul(y, w2, s2, b2, group_size=64, bits=4)
        return ops.argmax(x), x

    try:
        ops.set_pdl(False)
        ids0, x_off = chain()
        ops.set_pdl(True)
        ids1, x_on = chain()
    finally:
        ops.set_pdl(True)
    torch.cuda.synchronize()
    assert torch.equal(x_off, x_on)
    assert torch.equal(ids0, ids1)


def test_patched_model_wrapper_matches_direct_sdpa_forward():
    """``MHAPagedAttentionRuntime.patch_model`` wraps every ``self_attn``; calling the wrapper under a step context
    gives the same bits (and the same cache contents) as the direct ``sdpa_forward`` the runner uses."""
    import dataclasses

    from vllm_metal_b200.attention import context as ctx_mod
    from vllm_metal_b200.attention.impls.sdpa import sdpa_forward
    from vllm_metal_b200.attention.impls.sdpa_wrapper import SDPAPagedAttentionWrapper
    from vllm_metal_b200.attention.runtime.mha import MHAPagedAttentionRuntime
    from vllm_metal_b200.lm.decoder import PRESETS, Model

    args = dataclasses.replace(PRESETS["tiny"], sinks=True, sliding_window=16,
                               layer_types=["sliding_attention", "full_attention"])
    model = Model(args, torch.bfloat16, "cuda", seed=5)
    inner = [layer.self_attn for layer in model.layers]
    rt = MHAPagedAttentionRuntime(num_layers=args.num_layers, num_kv_heads=args.n_kv_heads, head_dim=args.head_dim,
                                  block_size=16, dtype=torch.bfloat16, device="cuda",
                                  sliding_window_per_layer=args.sliding_window_per_layer())
    rt.initialize(12)
    rt.warm_up()
    assert rt.patch_model(model) == args.num_layers
    assert all(isinstance(layer.self_attn, SDPAPagedAttentionWrapper) for layer in model.layers)
    ref = MHAPagedAttentionRuntime(num_layers=args.num_layers, num_kv_heads=args.n_kv_heads, head_dim=args.head_dim,
                                   block_size=16, dtype=torch.bfloat16, device="cuda",
                                   sliding_window_per_layer=args.sliding_window_per_layer())
    ref.initialize(12)
    g = torch.Generator(device="cuda").manual_seed(1)
    x = torch.randn((40, args.hidden), device="cuda", generator=g).bfloat16()
    ctx_mod.prepare_unified([([3, 7], 20), ([5], 4)], [([9, 2, 11], 38, 0)], 16)
    try:
        ctx = ctx_mod.get_context()
        for i, layer in enumerate(model.layers):
            got = layer.self_attn(x)
            want = sdpa_forward(inner[i], x, ctx, ref.kv_cache, i)
            assert torch.equal(got, want)
            assert torch.equal(rt.kv_cache.key_caches[i], ref.kv_cache.key_caches[i])
    finally:
        ctx_mod.clear_context()
