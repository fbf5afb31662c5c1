"""Parity of the tcgen05 tile kernel (csrc/paged_attention_prefill.cu) for varlen prefill / mixed batches.

Contract: vllm_metal/metal/kernels_v2/pagedattention_tiled.metal:114-525 (routing paged_ops.cpp:463-497) and its
tensor-unit variant pagedattention_nax.metal:406-449.  Checked against the oracle (the reference's own
ref_paged_attn math, tools/attention_bench_utils.py:11-62) at the reference's bf16/fp16 bars, and against the
per-row generic kernel with the reference's NAX-vs-tiled ULP metric (tools/nax_prefill_parity.py:53-67,181-192:
worst <= 2 ULP, mean <= 0.05 ULP of the output dtype).
"""

from __future__ import annotations

import numpy as np
import pytest
import torch

from .gpu_helpers import assert_close, make_paged_case, oracle_attention, run_attention, to_np

pytestmark = pytest.mark.gpu

GENERIC, TILE = 1, 3

SHAPES = [
    # (q_len, kv_len) per sequence
    [(128, 128)], [(129, 259)], [(300, 300), (17, 600)], [(256, 1300), (16, 16), (77, 1077)],
    [(129, 259), (1, 999), (64, 64), (1, 333)],  # tools/nax_prefill_parity.py:45-49
    [(512, 2048 + 512)],
]


_DTYPE_EPS = {"float16": 2.0 ** -10, "bfloat16": 2.0 ** -7}  # tools/nax_prefill_parity.py:65
MAX_ULP, MEAN_ULP = 2.0, 0.05                                # :66-67


def _ulp_metric(dtype: str, got: np.ndarray, ref: np.ndarray) -> tuple[float, float]:
    """The reference's divergence measure (tools/nax_prefill_parity.py:181-185): ULPs of the output dtype at the
    batch's peak magnitude."""
    delta = np.abs(got - ref)
    ulp = _DTYPE_EPS[dtype] * max(float(np.abs(ref).max()), 1.0)
    return float(delta.max()) / ulp, float(delta.mean()) / ulp


@pytest.mark.parametrize("seqs", SHAPES)
@pytest.mark.parametrize("hq,hkv,d", [(8, 2, 128), (8, 8, 64)])
@pytest.mark.parametrize("dtype", ["bfloat16", "float16"])
def test_tile_kernel_vs_oracle(ops, seqs, hq, hkv, d, dtype):
    c = make_paged_case(seqs, hq, hkv, d, 16, dtype, seed=len(seqs) + d)
    want = oracle_attention(c)
    got = run_attention(ops, c, kernel=TILE)
    assert_close(got, want, dtype, "tile kernel")
    # the reference's tensor-unit parity metric, against the per-row kernel
    row = to_np(run_attention(ops, c, kernel=GENERIC))
    mx_ulp, mean_ulp = _ulp_metric(dtype, to_np(got), row)
    print(f"tile vs per-row kernel: max {mx_ulp:.2f} ulp, mean {mean_ulp:.4f} ulp")
    assert mx_ulp <= MAX_ULP and mean_ulp <= MEAN_ULP, (mx_ulp, mean_ulp)


@pytest.mark.parametrize("window", [16, 100, 128, 1000])
@pytest.mark.parametrize("bs", [8, 16, 32])
def test_tile_kernel_window_and_block_sizes(ops, window, bs):
    c = make_paged_case([(200, 700), (130, 130), (1, 400)], 8, 4, 128, bs, "bfloat16", seed=window + bs)
    assert_close(run_attention(ops, c, window, kernel=TILE), oracle_attention(c, window), "bfloat16", f"W={window} bs={bs}")


def test_tile_kernel_sinks_gpt_oss_shape(ops):
    """gpt-oss attention: 64 q / 8 kv heads x 64, one sink logit per q head, window 128 (cfg d prefill)."""
    c = make_paged_case([(257, 257), (40, 900)], 64, 8, 64, 16, "bfloat16", seed=9, sinks=True)
    for w in (-1, 128):
        assert_close(run_attention(ops, c, w, kernel=TILE), oracle_attention(c, w), "bfloat16", f"sinks W={w}")


def test_mixed_batch_auto_route_splits_rows(ops):
    """Auto routing: long sequences on the tile kernel, decode rows on the per-row kernels, one call."""
    seqs = [(1, 4000), (1, 17), (2048, 2048 + 64), (1, 900), (5, 300), (333, 333)]
    c = make_paged_case(seqs, 32, 8, 128, 16, "bfloat16", seed=3)
    want = oracle_attention(c)
    out = torch.full_like(c["q"], 3.0)
    got = run_attention(ops, c, out=out)
    assert_close(got, want, "bfloat16", "mixed auto")


def test_garbage_beyond_seq_len_does_not_leak(ops):
    """Pages past seq_len hold NaN (never written): masked positions must not reach the output (0 * NaN)."""
    c = make_paged_case([(140, 140)], 8, 2, 128, 16, "bfloat16", seed=1)
    want = oracle_attention(c)
    kc, vc = c["kc"], c["vc"]
    bt = c["np"]["bt"]
    last = int(bt[0][140 // 16])
    kc[last, 140 % 16:] = float("nan")
    vc[last, 140 % 16:] = float("nan")
    kc[0] = float("nan")
    vc[0] = float("nan")
    assert_close(run_attention(ops, c, kernel=TILE), want, "bfloat16", "NaN pages")
