"""Allocation-free paged attention entry point (``b200_paged_attention_static_scratch``).

Not part of the reference's surface (include/b200_paged_ops.h says so): split-KV scratch in a file-scope device
array for callers that cannot call cudaMalloc.  Checked here through the C-ABI against the existing entry point
(``b200_paged_attention`` with a caller workspace) on two calls in a row -- the second with another shape, so the
first call's leftovers sit in the array -- and against the CPU oracle.  Collected last on purpose (pytest -x).
"""

from __future__ import annotations

import ctypes

import pytest
import torch

from tests.gpu_helpers import assert_close, make_paged_case, oracle_attention

pytestmark = pytest.mark.gpu

STATIC_SCRATCH_BYTES = 32 << 20  # B200_STATIC_SCRATCH_BYTES


@pytest.fixture(scope="module")
def ops():
    from vllm_metal_b200.metal import get_ops

    return get_ops()


def _args(c, out, max_seq_len, force_partitions, sliding_window=-1):
    """The struct ``paged_attention_primitive`` builds (metal/__init__.py), workspace left NULL."""
    from vllm_metal_b200.metal import PagedAttnArgs, _dtype_code

    a = PagedAttnArgs()
    q, kc = c["q"], c["kc"]
    a.out, a.query = out.data_ptr(), q.data_ptr()
    a.key_cache, a.value_cache = kc.data_ptr(), c["vc"].data_ptr()
    a.block_tables, a.seq_lens, a.cu_seqlens_q = c["bt"].data_ptr(), c["seq_lens"].data_ptr(), c["cu"].data_ptr()
    a.sinks = None if c["sinks"] is None else c["sinks"].data_ptr()
    a.workspace, a.workspace_bytes = None, 0
    a.num_q_tokens, a.num_seqs = q.shape[0], c["cu"].shape[0] - 1
    a.num_q_heads, a.num_kv_heads, a.head_dim = c["hq"], c["hkv"], c["d"]
    a.block_size = c["block_size"]
    a.max_blocks_per_seq = c["bt"].shape[1]
    a.max_seq_len = max_seq_len
    a.sliding_window = sliding_window
    a.dtype = _dtype_code("query", q)
    a.force_partitions = force_partitions
    a.kernel = 0
    a.q_stride, a.out_stride = q.stride(0), out.stride(0)
    a.kv_block_stride, a.kv_token_stride, a.kv_head_stride = kc.stride()[:3]
    a.num_blocks = kc.shape[0]
    a.scale, a.softcap = c["scale"], 0.0
    a.out_dtype = a.dtype
    a.window_seqlen_q = 1
    return a


def _both(ops, c, force_partitions, sliding_window=-1):
    lib = ops._lib
    stream = torch.cuda.current_stream().cuda_stream
    ref = torch.full_like(c["q"], 3.0)
    ops.paged_attention_primitive(
        c["q"], c["kc"], c["vc"], c["hkv"], c["scale"], 0.0, c["bt"], c["seq_lens"], c["cu"], c["block_size"],
        c["max_seq_len"], sliding_window, ref, sinks=c["sinks"], _force_partitions=force_partitions)
    got = torch.full_like(c["q"], 5.0)
    a = _args(c, got, c["max_seq_len"], force_partitions, sliding_window)
    need = lib.b200_paged_attention_workspace_bytes(ctypes.byref(a))
    ops._check(lib.b200_paged_attention_static_scratch(ctypes.byref(a), stream))
    torch.cuda.synchronize()
    return ref, got, need


def test_static_scratch_matches_the_workspace_entry_point_twice_in_a_row(ops):
    # call 1: decode rows, forced 512-token partitions (the scratch IS used), GQA 32q/8kv x 128, bf16
    c1 = make_paged_case([(1, 3000), (1, 517), (1, 2049), (1, 64)], 32, 8, 128, 16, dtype="bfloat16", seed=31)
    ref1, got1, need1 = _both(ops, c1, force_partitions=2)
    assert 0 < need1 <= STATIC_SCRATCH_BYTES
    assert torch.equal(got1, ref1)  # same plan, same kernels: the same bits
    assert_close(got1, oracle_attention(c1), "bfloat16", "static scratch, call 1")

    # call 2, straight after, another shape / dtype / sinks / window: call 1's partials are still in the array
    c2 = make_paged_case([(1, 1500), (1, 4100), (1, 9)], 64, 8, 64, 16, dtype="float16", seed=32, sinks=True)
    ref2, got2, need2 = _both(ops, c2, force_partitions=2, sliding_window=700)
    assert 0 < need2 <= STATIC_SCRATCH_BYTES
    assert torch.equal(got2, ref2)
    assert_close(got2, oracle_attention(c2, window=700), "float16", "static scratch, call 2")

    # and the planner's own choice (thin grid -> it splits by itself) plus a single-pass plan (no scratch at all)
    c3 = make_paged_case([(1, 8192)], 32, 8, 128, 16, dtype="bfloat16", seed=33)
    ref3, got3, _ = _both(ops, c3, force_partitions=0)
    assert torch.equal(got3, ref3)
    ref4, got4, need4 = _both(ops, c1, force_partitions=1)
    assert need4 == 0 and torch.equal(got4, ref4)


def test_static_scratch_rejects_plans_beyond_the_array(ops):
    """256 rows x 64 heads x 16 forced partitions x 130 floats = 136 MB > 32 MiB: refused before any launch."""
    c = make_paged_case([(1, 64)] * 256, 64, 8, 128, 16, dtype="bfloat16", seed=34)
    out = torch.zeros_like(c["q"])
    a = _args(c, out, 8192, force_partitions=2)  # max_seq_len drives the forced partition count
    assert ops._lib.b200_paged_attention_workspace_bytes(ctypes.byref(a)) > STATIC_SCRATCH_BYTES
    with pytest.raises(ValueError, match="B200_STATIC_SCRATCH_BYTES"):
        ops._check(ops._lib.b200_paged_attention_static_scratch(ctypes.byref(a),
                                                                torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    assert not out.any()
