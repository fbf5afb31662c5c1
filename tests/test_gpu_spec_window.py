"""Spec-decode window-mode paged attention parity (reference tests/test_spec_window_parity.py).

A K+1-row verification window can be dispatched expanded (K+1 one-row segments, ``window_seqlen_q=1``) or
windowed (one K+1-row segment with the hint), where the decode kernel packs floor(16 / G) query rows of the
sequence into one CTA and shares every KV read.  The reference pins the two BITWISE equal; so do these tests,
through the C-ABI, and both against the CPU oracle.
"""

from __future__ import annotations

import numpy as np
import pytest
import torch

from tests.gpu_helpers import assert_close, make_paged_case, oracle_attention, to_np

pytestmark = pytest.mark.gpu

GENERIC, MMA = 1, 2


@pytest.fixture(scope="module")
def ops():
    from vllm_metal_b200.metal import get_ops

    return get_ops()


def _run(ops, c, bt, seq_lens, cu, hint, window, max_kv, kernel=0, force_partitions=0):
    out = torch.full_like(c["q"], 7.0)
    ops.paged_attention_primitive(
        c["q"], c["kc"], c["vc"], c["hkv"], c["scale"], 0.0, bt, seq_lens, cu, c["block_size"], max_kv, window,
        out, window_seqlen_q=hint, sinks=c["sinks"], _kernel=kernel, _force_partitions=force_partitions)
    torch.cuda.synchronize()
    return out


def _expanded_and_windowed(ops, hq, hkv, dtype, ctx, windows, sliding_window, seed, d=128, block_size=16,
                           sinks=False, kernel=0, force_partitions=0):
    """Reference ``_expanded_and_windowed`` (tests/test_spec_window_parity.py:104-173): same query and cache, once
    as sum(windows) one-row segments with seq_len ctx+j+1, once as len(windows) segments with the hint."""
    c = make_paged_case([(w, ctx + w) for w in windows], hq, hkv, d, block_size, dtype=dtype, seed=seed,
                        sinks=sinks)
    max_kv = ctx + max(windows)
    rows = [s for s, w in enumerate(windows) for _ in range(w)]
    exp_bt = c["bt"][rows].contiguous()
    exp_sl = torch.tensor([ctx + j + 1 for w in windows for j in range(w)], dtype=torch.int32, device="cuda")
    exp_cu = torch.arange(len(rows) + 1, dtype=torch.int32, device="cuda")
    expanded = _run(ops, c, exp_bt, exp_sl, exp_cu, 1, sliding_window, max_kv, kernel, force_partitions)
    windowed = _run(ops, c, c["bt"], c["seq_lens"], c["cu"], max(windows), sliding_window, max_kv, kernel,
                    force_partitions)
    return expanded, windowed, c


@pytest.mark.parametrize("dtype", ["float16", "bfloat16", "float32"])
@pytest.mark.parametrize("hq,hkv", [(16, 8), (32, 8)])
@pytest.mark.parametrize("ctx", [512, 8192])
@pytest.mark.parametrize("window", [2, 4, 5, 6, 8])
def test_windowed_matches_expanded(ops, dtype, hq, hkv, ctx, window):
    """:176-194.  fp32 runs the scalar per-row kernel (the hint changes nothing there)."""
    expanded, windowed, _ = _expanded_and_windowed(ops, hq, hkv, dtype, ctx, [window], -1, seed=1)
    assert torch.equal(expanded, windowed)


@pytest.mark.parametrize("dtype", ["float16", "bfloat16"])
@pytest.mark.parametrize("hq,hkv", [(16, 8), (32, 8), (8, 8), (64, 8)])
@pytest.mark.parametrize("ctx", [512, 8192])
def test_windowed_matches_oracle(ops, dtype, hq, hkv, ctx):
    """:197-242, against the CPU oracle with the north-star tolerance."""
    _, windowed, c = _expanded_and_windowed(ops, hq, hkv, dtype, ctx, [5, 1, 3], -1, seed=2)
    assert_close(windowed, oracle_attention(c), dtype, f"window mode {hq}q/{hkv}kv ctx {ctx}")


@pytest.mark.parametrize("d", [64, 80, 96, 112, 192, 256])
def test_windowed_matches_expanded_head_sizes(ops, d):
    """:245-252 (64 runs the window-mode kernel, the others the per-row kernel)."""
    expanded, windowed, c = _expanded_and_windowed(ops, 16, 8, "float16", 1024, [6], -1, seed=4, d=d)
    assert torch.equal(expanded, windowed)
    assert_close(windowed, oracle_attention(c), "float16", f"head size {d}")


@pytest.mark.parametrize("block_size", [8, 32])
def test_windowed_matches_expanded_block_sizes(ops, block_size):
    """:255-261."""
    expanded, windowed, _ = _expanded_and_windowed(ops, 16, 8, "float16", 2048, [6], -1, seed=5,
                                                    block_size=block_size)
    assert torch.equal(expanded, windowed)


def test_windowed_matches_expanded_large_window(ops):
    """:264-274: windows past one CTA's rows run as several sub-windows."""
    for window in (12, 16, 23):
        expanded, windowed, c = _expanded_and_windowed(ops, 16, 8, "float16", 2048, [window], -1, seed=7)
        assert torch.equal(expanded, windowed), f"window={window}"
        assert_close(windowed, oracle_attention(c), "float16", f"window {window}")


def test_windowed_matches_expanded_zero_context_window(ops):
    """:277-288: a merged window over an empty prefix."""
    for window in (2, 4, 16):
        expanded, windowed, c = _expanded_and_windowed(ops, 16, 8, "float16", 0, [window], -1, seed=9)
        assert torch.equal(expanded, windowed), f"window={window}"
        assert_close(windowed, oracle_attention(c), "float16", f"zero-context window {window}")


def test_windowed_single_partition_family(ops):
    """:291-299: total KV inside one partition -> single pass."""
    assert 200 + 6 <= ops.PARTITION_SIZE
    expanded, windowed, _ = _expanded_and_windowed(ops, 16, 8, "float16", 200, [6], -1, seed=8,
                                                    force_partitions=1)
    assert torch.equal(expanded, windowed)


@pytest.mark.parametrize("hq,hkv,dtype", [(8, 8, "float16"), (32, 8, "bfloat16"), (16, 8, "float16")])
@pytest.mark.parametrize("ctx", [2048, 5000])
def test_windowed_forced_split_family(ops, hq, hkv, dtype, ctx):
    """:302-313: the partitioned family, forced (512-token partitions) -- rows of one CTA end in different
    partitions and the last partition holds only some of them."""
    expanded, windowed, c = _expanded_and_windowed(ops, hq, hkv, dtype, ctx, [6, 3], -1, seed=9,
                                                    force_partitions=2)
    assert torch.equal(expanded, windowed)
    assert_close(windowed, oracle_attention(c), dtype, "forced split")


def test_windowed_rows_straddle_a_partition_boundary(ops):
    """ctx chosen so the window's rows end either side of a 512-token partition edge."""
    for ctx in (509, 510, 511, 512, 1021):
        expanded, windowed, c = _expanded_and_windowed(ops, 32, 8, "bfloat16", ctx, [4], -1, seed=10,
                                                        force_partitions=2)
        assert torch.equal(expanded, windowed), f"ctx={ctx}"
        assert_close(windowed, oracle_attention(c), "bfloat16", f"ctx {ctx}")


@pytest.mark.parametrize(
    "cu,total_q,hint,match",
    [
        ([0, 6], 6, 2, r"window_seqlen_q=2 does not match the longest cu_seqlens_q segment \(6\)"),
        ([0, 1, 6], 6, 3, r"window_seqlen_q=3 does not match the longest cu_seqlens_q segment \(5\)"),
        ([0, 2, 4], 4, 4, r"window_seqlen_q=4 does not match the longest cu_seqlens_q segment \(2\)"),
        ([1, 4], 4, 3, r"cu_seqlens_q must start at 0 and end at the query row count \(4\)"),
        ([0, 3], 4, 3, r"cu_seqlens_q must start at 0 and end at the query row count \(4\)"),
        ([0, 0, 4], 4, 4, r"cu_seqlens_q segment 0 is empty \(0 -> 0\)"),
        ([0, 4], 4, 0, r"window_seqlen_q must be >= 1 \(1 = per-token decode\), got 0"),
        ([0, 4], 4, -4, r"window_seqlen_q must be >= 1 \(1 = per-token decode\), got -4"),
    ],
)
def test_window_hint_contract_rejected(ops, cu, total_q, hint, match):
    """:316-412: the hint must equal the longest segment of a structurally valid cu_seqlens_q."""
    g = torch.Generator().manual_seed(10)
    kc = torch.randn((4, 16, 8, 128), generator=g).half().cuda()
    vc = torch.randn((4, 16, 8, 128), generator=g).half().cuda()
    q = torch.randn((total_q, 16, 128), generator=g).half().cuda()
    nseg = len(cu) - 1
    bt = torch.arange(4, dtype=torch.int32).repeat(nseg, 1).cuda()
    with pytest.raises(ValueError, match=match):
        ops.paged_attention_primitive(
            q, kc, vc, 8, 128 ** -0.5, 0.0, bt, torch.full((nseg,), 56, dtype=torch.int32).cuda(),
            torch.tensor(cu, dtype=torch.int32).cuda(), 16, 56, -1, torch.zeros_like(q), window_seqlen_q=hint)


def test_window_hint_rejects_non_int32_cu(ops):
    """:415-443."""
    g = torch.Generator().manual_seed(13)
    kc = torch.randn((4, 16, 8, 128), generator=g).half().cuda()
    q = torch.randn((4, 16, 128), generator=g).half().cuda()
    with pytest.raises(ValueError, match="int32"):
        ops.paged_attention_primitive(
            q, kc, kc.clone(), 8, 128 ** -0.5, 0.0, torch.arange(4, dtype=torch.int32)[None].cuda(),
            torch.tensor([56], dtype=torch.int32).cuda(), torch.tensor([0, 4], dtype=torch.int64).cuda(), 16, 56,
            -1, torch.zeros_like(q), window_seqlen_q=4)


def test_window_hint_rejected_past_head_size_bound(ops):
    """:446-471."""
    g = torch.Generator().manual_seed(14)
    kc = torch.randn((4, 16, 2, 512), generator=g).half().cuda()
    q = torch.randn((4, 4, 512), generator=g).half().cuda()
    with pytest.raises(ValueError, match=r"window_seqlen_q=4 requires head_size <= 256, got 512"):
        ops.paged_attention_primitive(
            q, kc, kc.clone(), 2, 512 ** -0.5, 0.0, torch.arange(4, dtype=torch.int32)[None].cuda(),
            torch.tensor([56], dtype=torch.int32).cuda(), torch.tensor([0, 4], dtype=torch.int32).cuda(), 16, 56,
            -1, torch.zeros_like(q), window_seqlen_q=4)


@pytest.mark.parametrize("hq,hkv,dtype", [(16, 8, "float16"), (32, 8, "bfloat16")])
@pytest.mark.parametrize("poison", [float("inf"), float("nan")])
def test_windowed_skips_masked_inf_v(ops, hq, hkv, dtype, poison):
    """:474-518: a non-finite V at a position only later window rows may attend must not reach earlier rows
    (0 * inf in the shared P.V product)."""
    ctx, window = 48, 4
    c = make_paged_case([(window, ctx + window)], hq, hkv, 128, 16, dtype=dtype, seed=11)
    pos = ctx + 2  # visible to window rows 2 and 3 only
    blk = int(c["bt"][0, pos // 16])
    c["vc"][blk, pos % 16] = poison
    c["kc"][blk, pos % 16, :, 0] = poison  # and a non-finite K there must stay behind the mask too
    rows = [0] * window
    exp_bt = c["bt"][rows].contiguous()
    exp_sl = torch.tensor([ctx + j + 1 for j in range(window)], dtype=torch.int32, device="cuda")
    exp_cu = torch.arange(window + 1, dtype=torch.int32, device="cuda")
    expanded = _run(ops, c, exp_bt, exp_sl, exp_cu, 1, -1, ctx + window)
    windowed = _run(ops, c, c["bt"], c["seq_lens"], c["cu"], window, -1, ctx + window)
    assert torch.equal(expanded[:2], windowed[:2])
    assert bool(torch.isfinite(windowed[:2].float()).all())
    assert not bool(torch.isfinite(windowed[2:].float()).all())  # rows 2-3 legitimately read it


@pytest.mark.parametrize("sliding_window", [512, 100, 16, 5])
@pytest.mark.parametrize("window", [2, 6])
@pytest.mark.parametrize("fp", [0, 2])
def test_windowed_matches_expanded_sliding_window(ops, sliding_window, window, fp):
    """:521-530; rows of one CTA start their windows in different 16-token tiles and (forced split) partitions."""
    for ctx in (8192, 1000, 3):
        expanded, windowed, c = _expanded_and_windowed(ops, 16, 8, "float16", ctx, [window], sliding_window,
                                                        seed=2, force_partitions=fp)
        assert torch.equal(expanded, windowed), f"ctx={ctx}"
        assert_close(windowed, oracle_attention(c, window=sliding_window), "float16", f"W={sliding_window}")


def test_windowed_matches_expanded_mixed_batch(ops):
    """:533-539: a plain decode row and two verify windows of different lengths."""
    expanded, windowed, c = _expanded_and_windowed(ops, 16, 8, "float16", 8192, [1, 4, 6], -1, seed=3)
    assert torch.equal(expanded, windowed)
    assert_close(windowed, oracle_attention(c), "float16", "mixed windows")


def test_windowed_with_sinks(ops):
    """gpt-oss shape: 64q/8kv x 64, sinks, sliding window 128 (pagedattention.metal:1318-1328 in window mode)."""
    for sw in (-1, 128):
        expanded, windowed, c = _expanded_and_windowed(ops, 64, 8, "bfloat16", 4096, [2, 2, 1], sw, seed=6, d=64,
                                                        sinks=True)
        assert torch.equal(expanded, windowed)
        assert_close(windowed, oracle_attention(c, window=sw), "bfloat16", f"sinks W={sw}")


def test_windowed_leaves_rows_of_missing_subwindows_alone(ops):
    """Segments shorter than the hint own fewer sub-window CTAs; nothing outside [0, T) rows is written."""
    c = make_paged_case([(1, 700), (9, 709), (2, 64)], 32, 8, 128, 16, dtype="bfloat16", seed=12)
    out = _run(ops, c, c["bt"], c["seq_lens"], c["cu"], 9, -1, 709)
    assert_close(out, oracle_attention(c), "bfloat16", "ragged windows")


def test_window_mode_full_size_matches_expanded(ops):
    """cfg b's shape as a K+1 = 4 verification step: 64 sequences at ctx 4096, 32q/8kv x 128, 16-token pages,
    non-contiguous tables: bitwise equal to the expanded dispatch, oracle on a sample of sequences."""
    windows = [4] * 64
    expanded, windowed, c = _expanded_and_windowed(ops, 32, 8, "bfloat16", 4096, windows, -1, seed=21)
    assert torch.equal(expanded, windowed)
    n = c["np"]
    pick = [0, 17, 63]
    sub = dict(c)
    sub["np"] = dict(q=np.concatenate([n["q"][4 * s:4 * s + 4] for s in pick]), kc=n["kc"], vc=n["vc"],
                     bt=n["bt"][pick], seq_lens=n["seq_lens"][pick],
                     cu=np.arange(0, 4 * len(pick) + 1, 4).astype(np.int32), sinks=None)
    want = oracle_attention(sub)
    got = np.concatenate([to_np(windowed[4 * s:4 * s + 4]) for s in pick])
    assert_close(got, want, "bfloat16", "cfg-b verify window")
