"""Parity at the layer shapes BASELINE.json names (configs[2..4]), not just toy sizes.

* every linear of Qwen3-32B (int8 g64, K = 5120 / 8192 / 25600), Llama-3.1-70B (int4 g64,
  K = 8192 / 28672, and the TP=8 row-parallel shards K = 1024 / 3584) and gpt-oss-20b attention
  (int4 g64, K = 2880 = 45 groups) at its FULL size on the GPU; the oracle (fp32 ``x @ (s*q+b)^T``)
  is evaluated on a seeded subset of output rows so the CPU side stays in seconds;
* one Qwen3-32B-shaped layer and a gpt-oss-shaped sliding+full layer pair end to end
  (prefill chunk + decode rows, sinks, window 128, bs 32) against ``OracleDecoder``.

Tolerance: north-star bf16 bound, 2e-2 max-abs + 1e-2 relative.
"""

from __future__ import annotations

import dataclasses

import numpy as np
import pytest
import torch

from oracle import affine_quant as oq

from .gpu_helpers import assert_close, oracle_from_model, to_np

pytestmark = pytest.mark.gpu

SHAPES = [
    # name, N, K, bits, M
    ("qwen3-32b qkv", 10240, 5120, 8, 64), ("qwen3-32b o", 5120, 8192, 8, 64),
    ("qwen3-32b gate_up", 51200, 5120, 8, 33), ("qwen3-32b down", 5120, 25600, 8, 64),
    ("qwen3-32b gate_up tp4 M=2080", 12800, 5120, 8, 2080), ("qwen3-32b down tp4 M=2080", 5120, 6400, 8, 2080),
    ("llama-70b qkv", 10240, 8192, 4, 16), ("llama-70b o", 8192, 8192, 4, 16),
    ("llama-70b gate_up", 57344, 8192, 4, 16), ("llama-70b down K=28672", 8192, 28672, 4, 16),
    ("llama-70b o tp8", 8192, 1024, 4, 16), ("llama-70b down tp8", 8192, 3584, 4, 16),
    ("llama-70b qkv tp8", 1280, 8192, 4, 16), ("llama-70b gate_up tp8", 7168, 8192, 4, 16),
    ("gpt-oss qkv", 5120, 2880, 4, 32), ("gpt-oss o", 2880, 4096, 4, 32),
]


@pytest.mark.parametrize("name,n,k,bits,m", SHAPES, ids=[s[0].replace(" ", "_") for s in SHAPES])
def test_linear_at_baseline_shape(ops, name, n, k, bits, m):
    g = 64
    gen = torch.Generator(device="cuda").manual_seed(n + k)
    words = k * bits // 32
    w = torch.randint(-2**31, 2**31 - 1, (n, words), dtype=torch.int64, device="cuda", generator=gen).to(torch.int32)
    # MLX-affine triples of N(0, 0.02^2)-like weights: s ~ 6 sigma / (2^bits - 1), b ~ -s * 2^(bits-1)
    s = (torch.rand((n, k // g), device="cuda", generator=gen) * 0.5 + 0.75) * (0.12 / (2**bits - 1))
    b = -s * 2 ** (bits - 1) * (1.0 + 0.1 * torch.randn((n, k // g), device="cuda", generator=gen))
    s, b = s.bfloat16(), b.bfloat16()
    x = torch.randn((m, k), device="cuda", generator=gen).bfloat16()
    from vllm_metal_b200.quant.linear import QuantizedLinear

    lin = QuantizedLinear(k, n, False, g, bits, torch.bfloat16, "cuda")
    lin.weight.copy_(w)
    lin.scales.copy_(s)
    lin.biases.copy_(b)
    lin.prepack(keep_mlx_layout=True)
    got = lin(x)
    got_mlx_layout = ops.quantized_matmul(x, w, s, b, group_size=g, bits=bits)
    torch.cuda.synchronize()
    rows = np.sort(np.random.default_rng(k).choice(n, size=min(n, 384), replace=False))
    rows[-1] = n - 1  # always include the last row (ragged last tile)
    rt = torch.from_numpy(rows).cuda()
    mm = min(m, 96)
    want = oq.quantized_matmul(to_np(x[:mm]), w[rt].cpu().numpy().view(np.uint32), to_np(s[rt]), to_np(b[rt]), g, bits,
                               dtype="bfloat16")
    assert_close(got[:mm][:, rt], want, "bfloat16", f"{name} (streaming layout)")
    assert_close(got_mlx_layout[:mm][:, rt], want, "bfloat16", f"{name} (MLX layout)")
    if m > mm:  # the tail rows against the head's oracle would need more CPU; check them on the device
        assert torch.isfinite(got.float()).all()
        assert (got - got_mlx_layout).abs().max().item() <= 2e-2 + 1e-2 * got.abs().max().item()


def _layer_run(args, decode_lens, prefill, block_size=16, seed=0):
    from vllm_metal_b200.attention.caches.kv_cache import B200PagedKVCache
    from vllm_metal_b200.lm.decoder import Model
    from vllm_metal_b200.v1.model_runner import PagedModelRunner

    model = Model(args, torch.bfloat16, "cuda", seed=seed)
    lens = [n + 1 for n in decode_lens] + [q + off for q, off in prefill]
    per = [(n + block_size - 1) // block_size for n in lens]
    nb = 1 + sum(per)
    cache = B200PagedKVCache(args.num_layers, args.n_kv_heads, args.head_dim, nb, block_size, torch.bfloat16, "cuda",
                             sliding_window_per_layer=args.sliding_window_per_layer())
    runner = PagedModelRunner(model, cache, max_batch=len(lens), max_model_len=max(lens) + block_size,
                              use_cuda_graph=False)
    oracle = oracle_from_model(model, nb, block_size, c_attention=True)
    rng = np.random.default_rng(seed + 4)
    perm = (rng.permutation(nb - 1) + 1).tolist()
    tables, off = [], 0
    for n in per:
        tables.append(perm[off:off + n])
        off += n
    nd = len(decode_lens)
    # 1) fill the contexts: one packed prefill of every sequence's history
    hist = [(tables[i], decode_lens[i], 0) for i in range(nd)] + \
           [(tables[nd + j], o, 0) for j, (_, o) in enumerate(prefill) if o > 0]
    flat = rng.integers(0, args.vocab, sum(h[1] for h in hist)).tolist()
    if flat:
        runner.execute(flat, [], hist)
        oracle.forward(flat, [], hist)
    # 2) the mixed step under test: decode rows first, then the prefill chunks at their offsets
    dec = [(tables[i], decode_lens[i]) for i in range(nd)]
    pre = [(tables[nd + j], q, o) for j, (q, o) in enumerate(prefill)]
    ids = rng.integers(0, args.vocab, nd + sum(q for q, _ in prefill)).tolist()
    got = runner.execute(ids, dec, pre)
    ologits = oracle.forward(ids, dec, pre)
    want = oracle.greedy(ologits).tolist()
    bad = []
    for i, (g_, w_) in enumerate(zip(got, want, strict=True)):
        if g_ != w_:
            gap = float(ologits[i, w_] - ologits[i, g_])
            bad.append((i, g_, w_, gap))
            assert gap <= 2e-2 + 1e-2 * abs(float(ologits[i, w_])), f"row {i}: {g_} vs {w_}, oracle gap {gap}"
    return len(got), bad


def test_qwen3_32b_layer_mixed_prefill_decode_int8():
    """cfg c's layer: hidden 5120, 64q/8kv x128, inter 25600, q/k norm, int8 g64; a 192-token chunk
    continuing at offset 320 + 12 decode rows (ctx 40..700)."""
    from vllm_metal_b200.lm.decoder import PRESETS

    args = dataclasses.replace(PRESETS["qwen3-32b"], num_layers=1, vocab=2048)
    dl = np.random.default_rng(4).integers(40, 300, 10).tolist() + [700, 1300]
    n, bad = _layer_run(args, dl, [(192, 320)])
    assert n == 13 and len(bad) <= 1, bad


def test_gpt_oss_attention_layers_sinks_window_bs32():
    """cfg d: gpt-oss-20b attention blocks (hidden 2880, 64q/8kv x64, sinks, W=128 on even layers,
    full on odd), int4 g64 q/k/v/o, 32 decode rows at ctx 130..600."""
    from vllm_metal_b200.lm.decoder import PRESETS

    args = dataclasses.replace(PRESETS["gpt-oss-20b-attn"], num_layers=2, vocab=2048,
                               layer_types=["sliding_attention", "full_attention"])
    dl = np.random.default_rng(5).integers(130, 200, 26).tolist() + [600, 450, 300, 1030, 129, 128]
    n, bad = _layer_run(args, dl, [])
    assert n == 32 and len(bad) <= 1, bad


def test_llama_70b_tp8_shard_layer_decode_bs16():
    """cfg e's per-rank layer: 8 q heads / 1 kv head x128, hidden 8192, inter 28672/8, int4 g64, 16 decode
    rows at long-ish ctx (split-KV engaged).  world=1 model with the shard's head counts: the row-parallel
    K slices (1024 / 3584) and the single-kv-head attention grid are what TP=8 runs per GPU."""
    from vllm_metal_b200.lm.decoder import ModelArgs

    args = ModelArgs("llama-70b-tp8-shard", 1, 8192, 8, 1, 128, 3584, 2048)
    # o_proj then has K = 8*128 = 1024 and down K = 3584; the oracle holds the same weights.
    dl = np.random.default_rng(6).integers(20, 300, 12).tolist() + [2300, 1500, 1100, 700]
    n, bad = _layer_run(args, dl, [])
    assert n == 16 and len(bad) <= 1, bad
