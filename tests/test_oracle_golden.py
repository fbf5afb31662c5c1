"""Pin the CPU oracle against the reference's golden vectors (CPU only).

Sources of truth, in order:
* literal known answers copied as DATA from the reference's tests
  (tests/test_paged_attention.py:132-323, tests/test_block_size_translation.py:30-108,
  tests/test_attention_sdpa.py:630-685, tests/test_reshape_and_cache.py:30-70);
* fixtures produced by executing the reference's own Python
  (tests/golden/make_golden.py -> metadata_golden.json, attention_golden.npz).
"""

from __future__ import annotations

import numpy as np
import pytest

from oracle import affine_quant, kv_write, metadata, paged_attention
from oracle._numeric import bf16_round, round_to

# ---------------------------------------------------------------- metadata --

LITERAL_PREPARE = [
    # (decode, prefill, block_size, merge, expected fields)
    ([], [([10, 11], 5, 0)], 4, False,
     dict(slot_mapping=[40, 41, 42, 43, 44], block_tables=[[10, 11]], context_lens=[5],
          cu_seqlens=[0, 5], offsets=[0])),
    ([], [([10], 3, 0), ([20], 2, 0)], 4, False,
     dict(slot_mapping=[40, 41, 42, 80, 81], cu_seqlens=[0, 3, 5], block_tables=[[10], [20]],
          context_lens=[3, 2], offsets=[0, 0])),
    ([], [([5, 6], 5, 0)], 4, False,
     dict(cu_seqlens=[0, 5], slot_mapping=[20, 21, 22, 23, 24], block_tables=[[5, 6]],
          context_lens=[5])),
    ([], [([10, 11], 3, 4)], 4, False,
     dict(cu_seqlens=[0, 3], slot_mapping=[44, 45, 46], block_tables=[[10, 11]],
          context_lens=[7], offsets=[4])),
    ([([5, 6], 7)], [], 4, False,
     dict(slot_mapping=[27], context_lens=[8], offsets=[7], cu_seqlens=[0, 1])),
    ([([5, 6, 7], 7, 3)], [], 4, True,
     dict(slot_mapping=[27, 28, 29], block_tables=[[5, 6, 7]], context_lens=[10], offsets=[7],
          cu_seqlens=[0, 3], verify_window_q=3, num_decode_requests=1)),
    ([([5, 6, 7], 7, 3)], [], 4, False,
     dict(slot_mapping=[27, 28, 29], block_tables=[[5, 6, 7]] * 3, context_lens=[8, 9, 10],
          offsets=[7, 8, 9], cu_seqlens=[0, 1, 2, 3], verify_window_q=1, num_decode_requests=1)),
    ([([5, 6, 7], 7, 3)], [([10, 11], 5, 0)], 4, True,
     dict(slot_mapping=[27, 28, 29, 40, 41, 42, 43, 44], block_tables=[[5, 6, 7], [10, 11]],
          cu_seqlens=[0, 3, 8], context_lens=[10, 5], offsets=[7, 0], verify_window_q=1,
          num_decode_requests=1)),
    ([([5, 6], 7)], [([10, 11], 5, 0)], 4, False,
     dict(slot_mapping=[27, 40, 41, 42, 43, 44], cu_seqlens=[0, 1, 6], offsets=[7, 0],
          context_lens=[8, 5], block_tables=[[5, 6], [10, 11]])),
]


@pytest.mark.parametrize("case", LITERAL_PREPARE)
def test_oracle_prepare_unified_literals(case):
    dec, pre, bs, merge, want = case
    got = metadata.prepare_unified(dec, pre, bs, merge_verify_windows=merge)
    for k, v in want.items():
        assert got[k] == v, k


def test_oracle_prepare_grouped_literals():
    got = metadata.prepare_grouped([([[3, 4], [8]], 17, 1)], [([[5, 6], [9]], 2, 0)], (16, 32))
    g0, g1 = got["kv_groups"]
    assert g0["slot_mapping"] == [4 * 16 + 1, 5 * 16, 5 * 16 + 1]
    assert g1["slot_mapping"] == [8 * 32 + 17, 9 * 32, 9 * 32 + 1]
    assert g0["block_tables"] == [[3, 4], [5, 6]]
    assert g1["block_tables"] == [[8], [9]]
    assert got["slot_mapping"] == g0["slot_mapping"]
    got = metadata.prepare_grouped([([[3], [8, 9]], 17, 2)], [], (32, 16))
    assert got["context_lens"] == [18, 19]
    assert got["kv_groups"][0]["block_tables"] == [[3], [3]]
    assert got["kv_groups"][1]["block_tables"] == [[8, 9], [8, 9]]
    # tests/test_attention_sdpa.py:630-685 per-group routing
    got = metadata.prepare_grouped([([[3], [8, 9]], 17, 1)], [([[4], [10]], 2, 0)], (32, 16))
    assert got["kv_groups"][0]["slot_mapping"] == [3 * 32 + 17, 4 * 32, 4 * 32 + 1]
    assert got["kv_groups"][0]["block_tables"] == [[3], [4]]
    assert got["kv_groups"][1]["slot_mapping"] == [9 * 16 + 1, 10 * 16, 10 * 16 + 1]
    bt, kbs = metadata.build_block_tables(got["kv_groups"][1]["block_tables"], 16)
    assert bt.tolist() == [[8, 9], [10, 0]] and kbs == 16


def test_oracle_prepare_matches_reference_fixtures(golden_metadata):
    for case in golden_metadata["prepare"]:
        dec = [tuple(r) for r in case["decode"]]
        pre = [tuple(r) for r in case["prefill"]]
        got = metadata.prepare_grouped(dec, pre, tuple(case["block_sizes"]), case["merge"])
        want = case["out"]
        for k in ("slot_mapping", "block_tables", "context_lens", "cu_seqlens", "offsets",
                  "num_decode_requests", "verify_window_q"):
            assert got[k] == want[k], k
        assert got["kv_groups"] == want["kv_groups"]


def test_oracle_block_size_translation_literals():
    for bs in (32, 16, 8):
        assert metadata.pick_kernel_block_size(bs) == bs
    assert metadata.pick_kernel_block_size(544) == 32
    assert metadata.pick_kernel_block_size(48) == 16
    assert metadata.pick_kernel_block_size(24) == 8
    with pytest.raises(ValueError, match="not divisible"):
        metadata.pick_kernel_block_size(7)
    bt, kbs = metadata.build_block_tables([[0, 1], [2]], 16)
    assert kbs == 16 and bt.tolist() == [[0, 1], [2, 0]]
    bt, kbs = metadata.build_block_tables([[0], [1]], 544)
    assert kbs == 32 and bt[0].tolist() == list(range(17)) and bt[1].tolist() == list(range(17, 34))
    bt, _ = metadata.build_block_tables([[0, 2]], 544)
    assert bt[0].tolist() == list(range(17)) + list(range(34, 51))
    bt, _ = metadata.build_block_tables([[0, 1], [2]], 544)
    assert bt.shape == (2, 34)
    assert bt[1].tolist() == list(range(34, 51)) + list(range(17))
    assert metadata.build_block_tables([[0, 1, 2]], 544)[0].shape == (1, 51)
    bt, kbs = metadata.build_block_tables([], 16)
    assert bt.shape == (0, 0) and kbs == 16
    bt, kbs = metadata.build_block_tables([], 544)
    assert bt.shape == (0, 0) and kbs == 544


def test_oracle_block_tables_match_reference_fixtures(golden_metadata):
    for case in golden_metadata["build_block_tables"]:
        bt, kbs = metadata.build_block_tables(case["raw"], case["cache_block_size"])
        assert kbs == case["kernel_block_size"]
        assert bt.tolist() == case["tables"]
    for cbs, want in golden_metadata["pick_kernel_block_size"]:
        if want is None:
            with pytest.raises(ValueError):
                metadata.pick_kernel_block_size(cbs)
        else:
            assert metadata.pick_kernel_block_size(cbs) == want


# ------------------------------------------------------------------ KV write --

@pytest.mark.parametrize("dtype", [np.float16, np.uint16, np.float32])
@pytest.mark.parametrize("h,d", [(2, 64), (4, 128)])
def test_oracle_reshape_and_cache_equals_flat_scatter(dtype, h, d):
    rng = np.random.default_rng(0)
    nb, bs, t = 4, 16, 5
    key = rng.integers(0, 60000, (t, h, d)).astype(dtype)
    val = rng.integers(0, 60000, (t, h, d)).astype(dtype)
    slot = np.array([3, 20, 7, 40, 1], dtype=np.int64)
    kc = np.zeros((nb, bs, h, d), dtype)
    vc = np.zeros((nb, bs, h, d), dtype)
    kv_write.reshape_and_cache(key, val, kc, vc, slot)
    fk = np.zeros((nb * bs, h, d), dtype)
    fk[slot] = key
    fv = np.zeros((nb * bs, h, d), dtype)
    fv[slot] = val
    assert np.array_equal(kc.reshape(-1, h, d), fk)
    assert np.array_equal(vc.reshape(-1, h, d), fv)


def test_oracle_reshape_and_cache_skips_negative():
    rng = np.random.default_rng(1)
    key = rng.standard_normal((3, 4, 128)).astype(np.float16)
    val = rng.standard_normal((3, 4, 128)).astype(np.float16)
    kc = np.zeros((4, 16, 4, 128), np.float16)
    vc = np.zeros_like(kc)
    kv_write.reshape_and_cache(key, val, kc, vc, np.array([5, -1, 9], np.int64))
    ref = np.zeros((64, 4, 128), np.float16)
    ref[[5, 9]] = key[[0, 2]]
    assert np.array_equal(kc.reshape(64, 4, 128), ref)


# ----------------------------------------------------------------- attention --

CASES = ["decode_gqa", "decode_window", "mixed", "mixed_window", "softcap", "half"]


def _case(g, name):
    hq, hkv, d, bs, win = (int(v) for v in g[f"{name}.meta"])
    cap = float(g[f"{name}.softcap"][0])
    dt = "float16" if g[f"{name}.q"].dtype == np.float16 else "float32"
    return dict(
        q=g[f"{name}.q"].astype(np.float32), kc=g[f"{name}.kc"].astype(np.float32),
        vc=g[f"{name}.vc"].astype(np.float32), bt=g[f"{name}.bt"],
        q_lens=[int(v) for v in g[f"{name}.q_lens"]], kv_lens=[int(v) for v in g[f"{name}.kv_lens"]],
        scale=float(d) ** -0.5, window=None if win < 0 else win, cap=None if cap == 0 else cap,
        dtype=dt, ref=g[f"{name}.ref"].astype(np.float32))


@pytest.mark.parametrize("name", CASES)
def test_oracle_ref_paged_attn_matches_reference_output(golden_attention, name):
    """Our restatement of tools/attention_bench_utils.py:11-62 vs the reference's
    own function executed on the same inputs (fixture)."""
    c = _case(golden_attention, name)
    got = paged_attention.ref_paged_attn(c["q"], c["kc"], c["vc"], c["q_lens"], c["kv_lens"], c["bt"],
                                         c["scale"], c["window"], c["cap"], dtype=c["dtype"])
    tol = 2e-3 if c["dtype"] == "float16" else 2e-5
    np.testing.assert_allclose(got, c["ref"], atol=tol, rtol=tol)


@pytest.mark.parametrize("name", CASES)
def test_oracle_kernel_exact_within_reference_tolerance(golden_attention, name):
    """The kernel-exact restatement (log2 domain, +1e-6) stays inside the
    tolerance the reference accepts between kernel and ref_paged_attn
    (tests/test_primitive_and_donation.py:141-146: atol 1.5e-2, rtol 1e-2)."""
    c = _case(golden_attention, name)
    cu = np.concatenate([[0], np.cumsum(c["q_lens"])]).astype(np.int32)
    got = paged_attention.paged_attention_kernel_exact(
        c["q"], c["kc"], c["vc"], c["bt"], np.asarray(c["kv_lens"], np.int32), cu, c["scale"],
        -1 if c["window"] is None else c["window"], 0.0 if c["cap"] is None else c["cap"],
        dtype=c["dtype"])
    np.testing.assert_allclose(got, c["ref"], atol=1.5e-2, rtol=1e-2)


def test_oracle_sinks_equal_softmax_with_extra_column():
    rng = np.random.default_rng(3)
    hq, hkv, d, bs = 8, 2, 64, 16
    kc = rng.standard_normal((6, bs, hkv, d)).astype(np.float32)
    vc = rng.standard_normal((6, bs, hkv, d)).astype(np.float32)
    q = rng.standard_normal((2, hq, d)).astype(np.float32)
    bt = np.array([[1, 3, 5], [0, 2, 4]], np.int32)
    sinks = rng.standard_normal(hq).astype(np.float32)
    a = paged_attention.ref_paged_attn(q, kc, vc, [1, 1], [40, 33], bt, d ** -0.5, sinks=sinks)
    b = paged_attention.paged_attention_kernel_exact(
        q, kc, vc, bt, np.array([40, 33], np.int32), np.array([0, 1, 2], np.int32), d ** -0.5,
        sinks=sinks)
    np.testing.assert_allclose(a, b, atol=2e-5, rtol=2e-5)
    c = paged_attention.ref_paged_attn(q, kc, vc, [1, 1], [40, 33], bt, d ** -0.5)
    assert np.abs(a - c).max() > 1e-3  # the sink does change the result


def test_oracle_empty_rows_left_untouched():
    q = np.ones((1, 2, 8), np.float32)
    kc = np.ones((1, 8, 2, 8), np.float32)
    out = np.full((1, 2, 8), 7.0, np.float32)
    paged_attention.paged_attention_kernel_exact(
        q, kc, kc, np.zeros((1, 1), np.int32), np.array([0], np.int32), np.array([0, 1], np.int32),
        1.0, out=out)
    assert (out == 7.0).all()


# ---------------------------------------------------------------- quantised --

@pytest.mark.parametrize("bits", [4, 8])
def test_oracle_pack_unpack_roundtrip_and_bit_layout(bits):
    rng = np.random.default_rng(0)
    q = rng.integers(0, 1 << bits, (5, 64)).astype(np.uint8)
    w = affine_quant.pack(q, bits)
    assert w.dtype == np.uint32 and w.shape == (5, 64 * bits // 32)
    assert np.array_equal(affine_quant.unpack(w, bits), q)
    per = 32 // bits
    # element k sits at bit offset (k % per) * bits of word k // per, lowest bits first
    k = 11
    assert ((int(w[2, k // per]) >> ((k % per) * bits)) & ((1 << bits) - 1)) == int(q[2, k])


@pytest.mark.parametrize("bits,group", [(4, 32), (4, 64), (4, 128), (8, 32), (8, 64), (8, 128)])
def test_oracle_quantize_dequantize_error_bound(bits, group):
    rng = np.random.default_rng(1)
    w = (rng.standard_normal((16, 256)) * 0.02).astype(np.float32)
    wq, s, b = affine_quant.quantize(w, group, bits, "bfloat16")
    assert wq.shape == (16, 256 * bits // 32) and s.shape == (16, 256 // group) == b.shape
    assert np.array_equal(s, bf16_round(s)) and np.array_equal(b, bf16_round(b))
    deq = affine_quant.dequantize(wq, s, b, group, bits)
    step = np.repeat(np.abs(s), group, axis=1)
    brep = np.repeat(np.abs(b), group, axis=1)
    # the edge-anchored scale (s = edge / round(edge/s0)) can shrink the covered range by up
    # to half a step at the far end, so the bound is one step, plus the bf16 rounding
    # (2^-8 relative) of scale*q and bias
    bound = 1.0 * step + (1 << bits) * step * 2.0 ** -8 + brep * 2.0 ** -8 + 1e-6
    assert (np.abs(deq - w) <= bound).all()


def test_oracle_awq_repack_semantics():
    """tests/quant/test_awq_repack_parity.py:100-176: shapes + W' = (q - z) * s."""
    rng = np.random.default_rng(2)
    in_f, out_f, g = 256, 64, 128
    q = rng.integers(0, 16, (in_f, out_f)).astype(np.uint8)
    z = rng.integers(0, 16, (in_f // g, out_f)).astype(np.uint8)
    s = (rng.random((in_f // g, out_f)) * 0.01 + 0.001).astype(np.float16).astype(np.float32)
    qweight = affine_quant.awq_pack_int32(q)
    qzeros = affine_quant.awq_pack_int32(z)
    assert np.array_equal(affine_quant.awq_unpack_int32(qweight), q)
    w, sc, bi = affine_quant.awq_to_mlx_affine(qweight, qzeros, s, "float16")
    assert w.shape == (out_f, in_f // 8) and w.dtype == np.uint32
    assert sc.shape == (out_f, in_f // g) == bi.shape
    deq = affine_quant.dequantize(w, sc, bi, g, 4)
    want = ((q.astype(np.float32) - np.repeat(z, g, axis=0)) * np.repeat(s, g, axis=0)).T
    tol = np.maximum(np.abs(want) * 2.0 ** -9, 2e-3)
    assert (np.abs(deq - want) <= tol).all()


def test_oracle_quantized_matmul_equals_dense_and_factored():
    """tests/test_gguf_mlx_native.py:70-86: matmul vs dense dequant at max|ref|*4e-6."""
    rng = np.random.default_rng(4)
    w = (rng.standard_normal((48, 128)) * 0.05).astype(np.float32)
    x = rng.standard_normal((3, 128)).astype(np.float32)
    for bits in (4, 8):
        wq, s, b = affine_quant.quantize(w, 32, bits, "float32")
        y = affine_quant.quantized_matmul(x, wq, s, b, 32, bits)
        ref = x @ affine_quant.dequantize(wq, s, b, 32, bits).T
        assert np.abs(y - ref).max() <= np.abs(ref).max() * 4e-6 + 1e-7
        yf = affine_quant.quantized_matmul_factored(x, wq, s, b, 32, bits)
        np.testing.assert_allclose(yf, ref, atol=1e-4, rtol=1e-4)


def test_round_to_bf16_is_rne():
    x = np.array([1.0, 1.00390625, 1.01171875, -3.14159], np.float32)
    r = round_to(x, "bfloat16")
    assert r[0] == 1.0 and r[1] == 1.0 and r[2] == np.float32(1.015625)
    import torch

    assert np.array_equal(r, torch.from_numpy(x).to(torch.bfloat16).float().numpy())


# ------------------------------------------------------------ decoder oracle --

def _tiny_oracle(seed=0, qk_norm=True, sinks=False, window=-1):
    from oracle.decoder import OracleDecoder, QWeight

    rng = np.random.default_rng(seed)
    hidden, hq, hkv, d, inter, vocab, nl = 128, 4, 2, 32, 256, 96, 2

    def qw(n, k):
        w = round_to((rng.standard_normal((n, k)) * 0.05).astype(np.float32), "bfloat16")
        return QWeight(*affine_quant.quantize(w, 32, 4, "bfloat16"), 32, 4)

    layers = [dict(input_norm=np.ones(hidden, np.float32), post_norm=np.ones(hidden, np.float32),
                   qkv=qw((hq + 2 * hkv) * d, hidden), o=qw(hidden, hq * d),
                   q_norm=np.ones(d, np.float32) if qk_norm else None,
                   k_norm=np.ones(d, np.float32) if qk_norm else None,
                   gate_up=qw(2 * inter, hidden), down=qw(hidden, inter),
                   sinks=rng.standard_normal(hq).astype(np.float32) if sinks else None, window=window)
              for _ in range(nl)]
    w = dict(embed=qw(vocab, hidden), lm_head=None, norm=np.ones(hidden, np.float32), layers=layers)
    return OracleDecoder(w, hq, hkv, d, 1e-6, 10000.0, 8, 8, "float32"), vocab


@pytest.mark.parametrize("sinks,window", [(False, -1), (True, 5)])
def test_oracle_decoder_prefill_equals_prefill_plus_decode(sinks, window):
    """Chunked prefill + decode through the paged cache must reproduce a one-shot
    prefill (fp32 oracle): validates slot mapping, RoPE offsets, causal/window masks."""
    toks = [3, 17, 40, 8, 91, 5, 66, 12, 30, 77, 2]
    table = [5, 2, 7]
    a, _ = _tiny_oracle(1, sinks=sinks, window=window)
    one_shot = a.forward(toks, [], [(table, len(toks), 0)])
    b, _ = _tiny_oracle(1, sinks=sinks, window=window)
    b.forward(toks[:4], [], [(table, 4, 0)])
    b.forward(toks[4:10], [], [(table, 6, 4)])  # continuation chunk
    stepped = b.forward(toks[10:], [(table, 10)], [])
    np.testing.assert_allclose(stepped, one_shot, atol=2e-5, rtol=2e-5)
    assert OracleDecoder_greedy(stepped) == OracleDecoder_greedy(one_shot)


def OracleDecoder_greedy(logits):
    from oracle.decoder import OracleDecoder

    return OracleDecoder.greedy(logits).tolist()


# ------------------------------------------------- C port vs the NumPy oracle --

def test_c_port_matches_numpy_oracle():
    """oracle/csrc/oracle_cpu.c (the timed CPU baseline) against oracle/*.py."""
    from oracle import cport
    from oracle._numeric import bf16_bits, bf16_from_bits

    rng = np.random.default_rng(0)
    hq, hkv, d, bs, nb = 8, 2, 64, 16, 20
    kc = bf16_round(rng.standard_normal((nb, bs, hkv, d)).astype(np.float32))
    vc = bf16_round(rng.standard_normal((nb, bs, hkv, d)).astype(np.float32))
    q_lens, kv_lens = [1, 5, 1], [70, 33, 200]
    q = bf16_round(rng.standard_normal((sum(q_lens), hq, d)).astype(np.float32))
    bt = rng.integers(0, nb, (3, 13)).astype(np.int32)
    cu = np.array([0, 1, 6, 7], np.int32)
    sinks = rng.standard_normal(hq).astype(np.float32)
    for window, sk in ((-1, None), (24, None), (-1, sinks)):
        want = paged_attention.paged_attention_kernel_exact(
            q, kc, vc, bt, np.asarray(kv_lens, np.int32), cu, d ** -0.5, window, 0.0, sk, dtype="bfloat16")
        got = cport.paged_attention(bf16_bits(q), bf16_bits(kc), bf16_bits(vc), bt, kv_lens, cu, d ** -0.5,
                                    window, 0.0, sk)
        np.testing.assert_allclose(bf16_from_bits(got), want, atol=1e-2, rtol=1e-2)
    # KV write: bitwise
    key = rng.integers(0, 65535, (4, hkv, d)).astype(np.uint16)
    val = rng.integers(0, 65535, (4, hkv, d)).astype(np.uint16)
    slots = np.array([7, -1, 300, 19], np.int64)
    a_k, a_v = np.zeros((nb, bs, hkv, d), np.uint16), np.zeros((nb, bs, hkv, d), np.uint16)
    b_k, b_v = a_k.copy(), a_v.copy()
    kv_write.reshape_and_cache(key, val, a_k, a_v, slots)
    cport.reshape_and_cache(key, val, b_k, b_v, slots)
    assert np.array_equal(a_k, b_k) and np.array_equal(a_v, b_v)
    # quantised: dequantise bit-exact, matmul within bf16 rounding
    for bits, group in ((4, 64), (8, 32), (4, 128)):
        w = (rng.standard_normal((40, 256)) * 0.03).astype(np.float32)
        wq, s, b = affine_quant.quantize(w, group, bits, "bfloat16")
        deq = cport.dequantize(wq, bf16_bits(s), bf16_bits(b), group, bits)
        assert np.array_equal(bf16_from_bits(deq), affine_quant.dequantize(wq, s, b, group, bits, "bfloat16"))
        x = bf16_round(rng.standard_normal((5, 256)).astype(np.float32))
        y = cport.qmatmul(bf16_bits(x), wq, bf16_bits(s), bf16_bits(b), group, bits)
        want = affine_quant.quantized_matmul(x, wq, s, b, group, bits, dtype="bfloat16")
        np.testing.assert_allclose(bf16_from_bits(y), want, atol=1e-2, rtol=1e-2)
    assert cport.num_threads() >= 1
