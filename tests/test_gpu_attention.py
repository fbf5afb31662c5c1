"""GPU parity: paged attention + reshape_and_cache through the C-ABI vs the oracle.

Shapes and tolerances follow the reference's tests:
tests/test_primitive_and_donation.py:86-271 (decode / varlen / head sizes),
tests/test_split_kv_decode.py:134-252 (split-KV, windows, fully-masked partitions),
tests/test_paged_attention_sinks.py:184-253,374-389 (sinks + argument validation),
tests/test_reshape_and_cache.py:30-70 (bit-exact scatter).
"""

from __future__ import annotations

import numpy as np
import pytest
import torch

from oracle import kv_write
from oracle import paged_attention as oattn

from .gpu_helpers import TORCH_DT, assert_close, make_paged_case, oracle_attention, run_attention, to_np

pytestmark = pytest.mark.gpu

GENERIC, MMA = 1, 2


# ------------------------------------------------------------ reshape_and_cache

@pytest.mark.parametrize("dtype", ["float16", "bfloat16", "float32"])
@pytest.mark.parametrize("h,d", [(2, 64), (4, 128)])
def test_reshape_and_cache_matches_scatter(ops, dtype, h, d):
    tdt = TORCH_DT[dtype]
    nb, bs, t = 4, 16, 5
    g = torch.Generator().manual_seed(0)
    key = torch.randn((t, h, d), generator=g).to(tdt)
    val = torch.randn((t, h, d), generator=g).to(tdt)
    slot = torch.tensor([3, 20, 7, 40, 1], dtype=torch.int64)
    kc = torch.zeros((nb, bs, h, d), dtype=tdt, device="cuda")
    vc = torch.zeros_like(kc)
    nk, nv = ops.reshape_and_cache(key.cuda(), val.cuda(), kc, vc, slot.cuda())
    torch.cuda.synchronize()
    assert nk.data_ptr() == kc.data_ptr() and nv.data_ptr() == vc.data_ptr()  # aliased in place
    view = torch.int16 if tdt != torch.float32 else torch.int32
    okc = np.zeros((nb, bs, h, d), np.int64)
    ovc = np.zeros_like(okc)
    kv_write.reshape_and_cache(key.view(view).numpy().astype(np.int64), val.view(view).numpy().astype(np.int64),
                               okc, ovc, slot.numpy())
    assert np.array_equal(nk.cpu().view(view).numpy().astype(np.int64), okc)
    assert np.array_equal(nv.cpu().view(view).numpy().astype(np.int64), ovc)


def test_reshape_and_cache_skips_negative_and_out_of_range_slots(ops):
    h, d = 4, 128
    key = torch.randn(4, h, d).half().cuda()
    val = torch.randn(4, h, d).half().cuda()
    kc = torch.zeros((4, 16, h, d), dtype=torch.float16, device="cuda")
    vc = torch.zeros_like(kc)
    ops.reshape_and_cache(key, val, kc, vc, torch.tensor([5, -1, 9, 64], dtype=torch.int64).cuda())
    ref = torch.zeros(64, h, d, dtype=torch.float16, device="cuda")
    ref[[5, 9]] = key[[0, 2]]
    assert torch.equal(kc.reshape(64, h, d), ref)


def test_reshape_and_cache_full_size_roundtrip(ops):
    """BASELINE config b row shape (8 kv heads x 128, bf16), 64 decode tokens into a
    16k-block pool: every written row reads back bit-exact, nothing else changes."""
    h, d, bs, nb, t = 8, 128, 16, 16384, 64
    g = torch.Generator(device="cuda").manual_seed(1)
    key = torch.randn((t, h, d), generator=g, device="cuda").bfloat16()
    val = torch.randn((t, h, d), generator=g, device="cuda").bfloat16()
    kc = torch.zeros((nb, bs, h, d), dtype=torch.bfloat16, device="cuda")
    vc = torch.zeros_like(kc)
    slots = torch.randperm(nb * bs, generator=g, device="cuda")[:t].to(torch.int64)
    ops.reshape_and_cache(key, val, kc, vc, slots)
    assert torch.equal(kc.reshape(-1, h, d)[slots], key)
    assert torch.equal(vc.reshape(-1, h, d)[slots], val)
    assert int((kc.reshape(-1, h * d).abs().sum(1) != 0).sum()) == t


def test_reshape_and_cache_empty_and_validation(ops):
    kc = torch.zeros((2, 16, 2, 64), dtype=torch.bfloat16, device="cuda")
    e = torch.zeros((0, 2, 64), dtype=torch.bfloat16, device="cuda")
    ops.reshape_and_cache(e, e, kc, kc.clone(), torch.zeros(0, dtype=torch.int64, device="cuda"))
    k = torch.zeros((1, 2, 64), dtype=torch.float16, device="cuda")
    with pytest.raises(ValueError):
        ops.reshape_and_cache(k, k, kc, kc.clone(), torch.zeros(1, dtype=torch.int64, device="cuda"))
    kb = k.bfloat16()
    with pytest.raises(ValueError):
        ops.reshape_and_cache(kb, kb, kc, kc.clone(), torch.zeros(1, dtype=torch.int32, device="cuda"))


def test_copy_blocks(ops):
    kc = torch.randn((8, 16, 2, 64), device="cuda").bfloat16()
    vc = torch.randn((8, 16, 2, 64), device="cuda").bfloat16()
    k0, v0 = kc.clone(), vc.clone()
    ops.copy_blocks(kc, vc, torch.tensor([1, 3]), torch.tensor([5, 6]))
    k0[[5, 6]] = k0[[1, 3]]
    v0[[5, 6]] = v0[[1, 3]]
    assert torch.equal(kc, k0) and torch.equal(vc, v0)


def test_cache_copy_blocks_gathers_before_it_scatters():
    """``array[dst] = array[src]`` (reference kv_cache.py:353) reads every source before any write: a chain
    1 -> 2, 2 -> 3 must put the OLD block 2 into block 3."""
    from vllm_metal_b200.attention.caches.kv_cache import B200PagedKVCache

    c = B200PagedKVCache(2, 2, 64, 8, 16, torch.bfloat16, "cuda")
    for t in (*c.key_caches, *c.value_caches):
        t.copy_(torch.randn(t.shape, device="cuda").bfloat16())
    k0 = [k.clone() for k in c.key_caches]
    v0 = [v.clone() for v in c.value_caches]
    c.copy_blocks([(1, 2), (2, 3), (5, 6)])
    for k, v, kk, vv in zip(c.key_caches, c.value_caches, k0, v0, strict=True):
        kk[[2, 3, 6]] = kk[[1, 2, 5]]  # advanced indexing gathers first, like mx
        vv[[2, 3, 6]] = vv[[1, 2, 5]]
        assert torch.equal(k, kk) and torch.equal(v, vv)
    with pytest.raises(RuntimeError, match="twice"):
        c.copy_blocks([(1, 4), (2, 4)])


# ------------------------------------------------------------ golden fixtures

@pytest.mark.parametrize("name", ["decode_gqa", "decode_window", "mixed", "mixed_window", "softcap", "half"])
@pytest.mark.parametrize("kernel", [0, GENERIC])
def test_attention_matches_reference_golden(ops, golden_attention, name, kernel):
    """Outputs of the reference's own ref_paged_attn (fixtures), reference tolerance
    atol 1.5e-2 / rtol 1e-2 (tests/test_primitive_and_donation.py:141-146)."""
    g = golden_attention
    hq, hkv, d, bs, win = (int(v) for v in g[f"{name}.meta"])
    cap = float(g[f"{name}.softcap"][0])
    tdt = torch.float16 if g[f"{name}.q"].dtype == np.float16 else torch.float32
    q = torch.from_numpy(g[f"{name}.q"]).to(tdt).cuda()
    kc = torch.from_numpy(g[f"{name}.kc"]).to(tdt).cuda()
    vc = torch.from_numpy(g[f"{name}.vc"]).to(tdt).cuda()
    bt = torch.from_numpy(g[f"{name}.bt"]).cuda()
    q_lens, kv_lens = g[f"{name}.q_lens"], g[f"{name}.kv_lens"]
    cu = torch.from_numpy(np.concatenate([[0], np.cumsum(q_lens)]).astype(np.int32)).cuda()
    sl = torch.from_numpy(kv_lens.astype(np.int32)).cuda()
    out = torch.zeros_like(q)
    ops.paged_attention_primitive(q, kc, vc, hkv, float(d) ** -0.5, cap, bt, sl, cu, bs,
                                  int(kv_lens.max()), win, out, _kernel=kernel)
    torch.cuda.synchronize()
    np.testing.assert_allclose(to_np(out), g[f"{name}.ref"].astype(np.float32), atol=1.5e-2, rtol=1e-2)


# ------------------------------------------------------------------- decode

@pytest.mark.parametrize("hq,hkv", [(4, 4), (8, 2), (32, 8), (64, 8)])
@pytest.mark.parametrize("d", [64, 128])
@pytest.mark.parametrize("window", [-1, 128])
@pytest.mark.parametrize("dtype", ["bfloat16", "float16"])
def test_decode_vs_oracle(ops, hq, hkv, d, window, dtype):
    seqs = [(1, kv) for kv in (1, 15, 16, 17, 130, 523, 2048)]
    c = make_paged_case(seqs, hq, hkv, d, 16, dtype, seed=hq * 7 + d)
    want = oracle_attention(c, window)
    for kernel in (MMA, GENERIC):
        for fp in (1, 2):
            got = run_attention(ops, c, window, force_partitions=fp, kernel=kernel)
            assert_close(got, want, dtype, f"kernel={kernel} parts={fp}")


@pytest.mark.parametrize("block_size", [8, 16, 32])
@pytest.mark.parametrize("dtype", ["bfloat16", "float32"])
def test_decode_block_sizes_and_fp32(ops, block_size, dtype):
    c = make_paged_case([(1, 77), (1, 300), (1, 1025)], 8, 2, 128, block_size, dtype, seed=3)
    want = oracle_attention(c)
    for fp in (0, 1, 2):
        assert_close(run_attention(ops, c, force_partitions=fp), want, dtype, f"parts={fp}")


@pytest.mark.parametrize("d", [80, 96, 112, 192, 256])
def test_generic_kernel_other_head_sizes(ops, d):
    """Head sizes the reference instantiates (pagedattention.metal:2228-2339) beyond 64/128."""
    c = make_paged_case([(1, 200), (3, 90)], 4, 2, d, 16, "float16", seed=d)
    assert_close(run_attention(ops, c), oracle_attention(c), "float16", f"d={d}")


@pytest.mark.parametrize("kv_lens", [[8192], [100, 2048], [700, 4096], [1300]])
@pytest.mark.parametrize("dtype", ["float16", "bfloat16", "float32"])
def test_split_kv_decode(ops, kv_lens, dtype):
    """tests/test_split_kv_decode.py:134-153 (Qwen3-0.6B head shape 16q/8kv/128)."""
    c = make_paged_case([(1, kv) for kv in kv_lens], 16, 8, 128, 16, dtype, seed=len(kv_lens))
    want = oracle_attention(c)
    single = run_attention(ops, c, force_partitions=1)
    split = run_attention(ops, c, force_partitions=2)
    auto = run_attention(ops, c)
    for got, what in ((single, "single"), (split, "split"), (auto, "auto")):
        assert_close(got, want, dtype, what)


@pytest.mark.parametrize("window", [1, 16, 128, 512, 700])
def test_split_kv_sliding_window_masked_partitions(ops, window):
    """tests/test_split_kv_decode.py:190-252: partitions wholly outside the window
    must contribute nothing and the result must be finite."""
    c = make_paged_case([(1, 4096), (1, 513), (1, 100)], 16, 8, 128, 16, "bfloat16", seed=11)
    want = oracle_attention(c, window)
    for kernel in (MMA, GENERIC):
        for fp in (1, 2):
            got = run_attention(ops, c, window, force_partitions=fp, kernel=kernel)
            assert_close(got, want, "bfloat16", f"W={window} kernel={kernel} parts={fp}")


def test_window_zero_and_empty_rows(ops):
    """W == 0 masks everything: out = 0/(0+1e-6) = 0 (kernel-exact), rows with no KV
    keep their previous contents (pagedattention.metal:917-920)."""
    c = make_paged_case([(1, 40), (1, 0)], 8, 2, 64, 16, "bfloat16", seed=5)
    # second sequence has no KV at all -> its row must stay untouched
    out = torch.full_like(c["q"], 7.0)
    init = np.full(c["np"]["q"].shape, 7.0, np.float32)
    want = oracle_attention(c, -1, out_init=init)
    got = run_attention(ops, c, out=out)
    assert_close(got, want, "bfloat16", "empty row")
    assert (to_np(got)[1] == 7.0).all()
    z = run_attention(ops, c, window=0)
    assert (to_np(z)[0] == 0).all()


# --------------------------------------------------------------- varlen / mixed

VARLEN = [
    [(1, 1328), (5, 18), (129, 463)],
    [(8, 128), (16, 256)],
    [(32, 512), (64, 1024)],
    [(129, 259), (1, 999), (64, 64), (1, 333)],  # tools/nax_prefill_parity.py:45-49 (q, q+ctx)
]


@pytest.mark.parametrize("seqs", VARLEN)
@pytest.mark.parametrize("window", [-1, 100])
def test_varlen_mixed_vs_oracle(ops, seqs, window):
    c = make_paged_case(seqs, 32, 8, 128, 16, "bfloat16", seed=len(seqs))
    want = oracle_attention(c, window)
    assert_close(run_attention(ops, c, window), want, "bfloat16", "auto")
    assert_close(run_attention(ops, c, window, kernel=GENERIC), want, "bfloat16", "generic")


def test_prefill_only_head_sizes(ops):
    for d in (64, 96, 128):
        c = make_paged_case([(70, 70), (33, 33)], 8, 4, d, 16, "float16", seed=d)
        assert_close(run_attention(ops, c), oracle_attention(c), "float16", f"d={d}")


# ---------------------------------------------------------------------- sinks

@pytest.mark.parametrize("hq,hkv,d", [(8, 8, 64), (8, 2, 128), (64, 8, 64)])
@pytest.mark.parametrize("kv", [37, 600, 2100])
def test_sinks_decode(ops, hq, hkv, d, kv):
    """tests/test_paged_attention_sinks.py:184-237; the oracle is softmax over
    [scores, sink] with the sink column dropped."""
    c = make_paged_case([(1, kv), (1, 5)], hq, hkv, d, 16, "bfloat16", seed=kv, sinks=True)
    n = c["np"]
    ref = oattn.ref_paged_attn(n["q"], n["kc"], n["vc"], [1, 1], [kv, 5], n["bt"], c["scale"],
                               sinks=n["sinks"], dtype="float32")
    want = oracle_attention(c)
    np.testing.assert_allclose(want, ref, atol=2e-2, rtol=2e-2)
    for kernel in (MMA, GENERIC):
        for fp in (1, 2):
            got = run_attention(ops, c, force_partitions=fp, kernel=kernel)
            assert_close(got, want, "bfloat16", f"kernel={kernel} parts={fp}")


def test_sink_not_multiplied_by_warp_count(ops):
    """Regression pinned by tests/test_paged_attention_sinks.py:227-237: one KV token,
    q.k = 0, sink = 0 -> out must be v/2 (a sink folded once per warp gives v/(1+W))."""
    d = 64
    q = torch.zeros((1, 4, d), dtype=torch.float32, device="cuda")
    kc = torch.zeros((2, 16, 4, d), dtype=torch.float32, device="cuda")
    vc = torch.ones_like(kc)
    out = torch.zeros_like(q)
    bt = torch.tensor([[1]], dtype=torch.int32, device="cuda")
    ops.paged_attention_primitive(q, kc, vc, 4, 1.0, 0.0, bt, torch.tensor([1], dtype=torch.int32, device="cuda"),
                                  torch.tensor([0, 1], dtype=torch.int32, device="cuda"), 16, 1, -1, out,
                                  sinks=torch.zeros(4, device="cuda"))
    assert (out - 0.5).abs().max().item() < 1e-4


def test_sinks_gpt_oss_prefill_shape(ops):
    """tests/test_paged_attention_sinks.py:240-253: 64q/8kv/hd64 bf16, window 128."""
    c = make_paged_case([(200, 200), (1, 777), (40, 300)], 64, 8, 64, 16, "bfloat16", seed=9, sinks=True)
    for window in (-1, 128):
        assert_close(run_attention(ops, c, window), oracle_attention(c, window), "bfloat16", f"W={window}")


def test_sinks_validation(ops):
    """Messages matched by tests/test_paged_attention_sinks.py:374-389."""
    c = make_paged_case([(1, 20)], 8, 2, 64, 16, "bfloat16")
    args = (c["q"], c["kc"], c["vc"], 2, c["scale"], 0.0, c["bt"], c["seq_lens"], c["cu"], 16, 20, -1,
            torch.zeros_like(c["q"]))
    with pytest.raises(ValueError, match="one entry per query head"):
        ops.paged_attention_primitive(*args, sinks=torch.zeros(7, device="cuda"))
    with pytest.raises(ValueError, match="1-D"):
        ops.paged_attention_primitive(*args, sinks=torch.zeros(8, 1, device="cuda"))
    with pytest.raises(ValueError, match="float32"):
        ops.paged_attention_primitive(*args, sinks=torch.zeros(8, device="cuda", dtype=torch.bfloat16))
    with pytest.raises(ValueError, match="TurboQuant"):
        ops.paged_attention_primitive(*args, use_turboquant=True, sinks=torch.zeros(8, device="cuda"))
    with pytest.raises(ValueError, match="window_seqlen_q"):
        ops.paged_attention_primitive(*args, window_seqlen_q=0)


# ------------------------------------------------------ full-size properties

def test_full_size_decode_properties(ops):
    """BASELINE config b shape (bs 64, ctx 4096, 32q/8kv/128, block 16) — too big for the
    NumPy oracle in seconds, so check (1) 4 sampled sequences against the oracle,
    (2) split-KV == single pass, (3) MMA kernel == generic kernel, (4) linearity in V."""
    seqs = [(1, 4096 + (i % 7)) for i in range(64)]
    c = make_paged_case(seqs, 32, 8, 128, 16, "bfloat16", seed=42)
    auto = run_attention(ops, c)
    single = run_attention(ops, c, force_partitions=1)
    generic = run_attention(ops, c, kernel=GENERIC, force_partitions=2)
    assert (auto.float() - single.float()).abs().max().item() <= 1e-2
    assert (auto.float() - generic.float()).abs().max().item() <= 1e-2
    # sampled oracle rows
    for s in (0, 17, 40, 63):
        sub = dict(c)
        n = c["np"]
        sub["np"] = dict(q=n["q"][s:s + 1], kc=n["kc"], vc=n["vc"], bt=n["bt"][s:s + 1],
                         seq_lens=n["seq_lens"][s:s + 1], cu=np.array([0, 1], np.int32), sinks=None)
        assert_close(auto[s:s + 1], oracle_attention(sub), "bfloat16", f"seq {s}")
    # linearity: attention(V * 2) == 2 * attention(V) exactly in bf16 (power-of-two scale)
    c2 = dict(c)
    c2["vc"] = c["vc"] * 2
    twice = run_attention(ops, c2)
    assert torch.equal(twice.float(), auto.float() * 2)


@pytest.mark.parametrize("kernel,fp", [(0, 0), (GENERIC, 1), (MMA, 1), (MMA, 2), (3, 0)])
def test_fp16_output_of_a_bf16_call_is_the_same_values(ops, kernel, fp):
    """args.out_dtype = fp16 on a bf16 call (every kernel: generic, tensor-core decode single pass / split +
    merge, tcgen05 tile): the bf16-rounded result stored as fp16 -- the bf16 output's values exactly, except that
    magnitudes below fp16's smallest subnormal (2^-24) round to it or to zero."""
    seqs = [(1, 700), (1, 33)] if kernel != 3 else [(140, 300), (20, 20)]
    c = make_paged_case(seqs, 8, 2, 128, 16, "bfloat16", seed=11, sinks=True)
    ref = run_attention(ops, c, force_partitions=fp, kernel=kernel)
    out16 = torch.zeros(c["q"].shape, dtype=torch.float16, device="cuda")
    got = run_attention(ops, c, force_partitions=fp, kernel=kernel, out=out16)
    assert got.dtype == torch.float16
    diff = (got.float() - ref.float()).abs()
    assert diff.max().item() <= 2.0 ** -24
    assert torch.equal(got.float()[ref.float().abs() >= 2.0 ** -14], ref.float()[ref.float().abs() >= 2.0 ** -14])
