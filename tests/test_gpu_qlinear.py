"""GPU parity: MLX-affine quantized linears + decoder neighbours vs the oracle.

Bit-exact: int4/int8 unpack -> dequantised values (``dequantize``).  Within the
north-star tolerance (2e-2 max-abs / 1e-2 relative, bf16): quantized matmul.
Reference anchors: gguf/mlx_native.py:212-239, tests/test_gguf_mlx_native.py:70-86.
"""

from __future__ import annotations

import numpy as np
import pytest
import torch

from oracle import affine_quant as oq
from oracle._numeric import round_to

from .gpu_helpers import TORCH_DT, assert_close, to_np

pytestmark = pytest.mark.gpu


def _quantized(n, k, bits, group, dtype="bfloat16", seed=0, std=0.02):
    rng = np.random.default_rng(seed)
    w = round_to((rng.standard_normal((n, k)) * std).astype(np.float32), dtype)
    wq, s, b = oq.quantize(w, group, bits, dtype)
    tdt = TORCH_DT[dtype]
    return (wq, s, b,
            torch.from_numpy(wq.view(np.int32)).cuda(), torch.from_numpy(s).to(tdt).cuda(),
            torch.from_numpy(b).to(tdt).cuda())


@pytest.mark.parametrize("bits", [4, 8])
@pytest.mark.parametrize("group", [32, 64, 128])
@pytest.mark.parametrize("dtype", ["bfloat16", "float16"])
def test_dequantize_bit_exact(ops, bits, group, dtype):
    wq, s, b, tw, ts, tb = _quantized(96, 512, bits, group, dtype, seed=bits + group)
    got = ops.dequantize(tw, ts, tb, group, bits)
    want = oq.dequantize(wq, s, b, group, bits, dtype)
    assert np.array_equal(to_np(got), want)
    ids = torch.tensor([[5, 0], [95, 5]], dtype=torch.int64, device="cuda")
    rows = ops.dequantize(tw, ts, tb, group, bits, ids=ids)
    assert rows.shape == (2, 2, 512)
    assert np.array_equal(to_np(rows), want[ids.cpu().numpy()])
    # fp32 output of the same words is the unrounded s*q+b
    f32 = ops.dequantize(tw, ts, tb, group, bits, dtype=torch.float32)
    assert np.array_equal(to_np(f32), oq.dequantize(wq, s, b, group, bits, "float32"))


@pytest.fixture(params=[1, 2], ids=["mma_sync", "tcgen05"])
def kernel(request, ops):
    ops.set_qlinear_kernel(request.param)
    yield request.param
    ops.set_qlinear_kernel(0)


@pytest.mark.parametrize("m", [1, 3, 16, 17, 64, 100, 256, 300, 513])
@pytest.mark.parametrize("bits,group", [(4, 64), (4, 32), (4, 128), (8, 64), (8, 32), (8, 128)])
def test_quantized_matmul_vs_oracle(ops, kernel, m, bits, group):
    n, k = 320, 1024
    wq, s, b, tw, ts, tb = _quantized(n, k, bits, group, seed=m)
    rng = np.random.default_rng(m + 1)
    x = round_to(rng.standard_normal((m, k)).astype(np.float32), "bfloat16")
    tx = torch.from_numpy(x).bfloat16().cuda()
    got = ops.quantized_matmul(tx, tw, ts, tb, transpose=True, group_size=group, bits=bits)
    want = oq.quantized_matmul(x, wq, s, b, group, bits, dtype="bfloat16")
    assert got.shape == (m, n) and got.dtype == torch.bfloat16
    assert_close(got, want, "bfloat16", f"M={m}")


@pytest.mark.parametrize("n,k", [(4096, 4096), (1024, 4096), (6144, 4096), (4096, 14336), (2880, 2880), (136, 192)])
def test_quantized_matmul_model_shapes(ops, kernel, n, k):
    """Llama-8B projections (q/o, k/v, fused qkv, down) and gpt-oss K=2880=45*64."""
    wq, s, b, tw, ts, tb = _quantized(n, k, 4, 64, seed=n % 97)
    x = round_to(np.random.default_rng(5).standard_normal((64, k)).astype(np.float32), "bfloat16")
    got = ops.quantized_matmul(torch.from_numpy(x).bfloat16().cuda(), tw, ts, tb, group_size=64, bits=4)
    want = oq.quantized_matmul(x, wq, s, b, 64, 4, dtype="bfloat16")
    assert_close(got, want, "bfloat16", f"{n}x{k}")


@pytest.mark.parametrize("m", [1, 8, 16, 33, 64])
@pytest.mark.parametrize("n,k,bits,group", [(256, 4096, 4, 64), (128, 2880, 4, 32), (512, 8192, 8, 128),
                                           (1000, 2048, 4, 128), (384, 1536, 8, 64)])
def test_tcgen05_cluster_split_k(ops, m, n, k, bits, group):
    """Few row tiles -> K is split over a thread-block cluster and reduced through DSMEM;
    includes K with a ragged last packed stage (2880 = 11.25 x 256) and a dense bias."""
    ops.set_qlinear_kernel(2)
    try:
        wq, s, b, tw, ts, tb = _quantized(n, k, bits, group, seed=n + m)
        x = round_to(np.random.default_rng(m).standard_normal((m, k)).astype(np.float32), "bfloat16")
        bias = round_to(np.random.default_rng(7).standard_normal(n).astype(np.float32), "bfloat16")
        got = ops.quantized_matmul(torch.from_numpy(x).bfloat16().cuda(), tw, ts, tb, group_size=group, bits=bits,
                                   bias=torch.from_numpy(bias).bfloat16().cuda())
        want = oq.quantized_matmul(x, wq, s, b, group, bits, bias=bias, dtype="bfloat16")
        assert_close(got, want, "bfloat16", f"M={m} {n}x{k}")
    finally:
        ops.set_qlinear_kernel(0)


def test_tcgen05_strided_rows(ops):
    """x and y as column slices of wider buffers (row strides != K, N)."""
    ops.set_qlinear_kernel(2)
    try:
        wq, s, b, tw, ts, tb = _quantized(192, 512, 4, 64, seed=9)
        xw = torch.randn((20, 1024), device="cuda").bfloat16()
        x = xw[:, 256:768]
        yw = torch.zeros((20, 400), dtype=torch.bfloat16, device="cuda")
        ops.quantized_matmul(x, tw, ts, tb, group_size=64, bits=4, out=yw[:, 8:200])
        want = oq.quantized_matmul(to_np(x), wq, s, b, 64, 4, dtype="bfloat16")
        assert_close(yw[:, 8:200], want, "bfloat16", "strided")
        assert (yw[:, :8] == 0).all() and (yw[:, 200:] == 0).all()
    finally:
        ops.set_qlinear_kernel(0)


def test_quantized_matmul_bias_fp16_and_leading_dims(ops):
    wq, s, b, tw, ts, tb = _quantized(128, 256, 4, 32, "float16", seed=2)
    x = round_to(np.random.default_rng(3).standard_normal((2, 5, 256)).astype(np.float32), "float16")
    bias = round_to(np.random.default_rng(4).standard_normal(128).astype(np.float32), "float16")
    got = ops.quantized_matmul(torch.from_numpy(x).half().cuda(), tw, ts, tb, group_size=32, bits=4,
                               bias=torch.from_numpy(bias).half().cuda())
    want = oq.quantized_matmul(x.reshape(10, 256), wq, s, b, 32, 4, bias=bias, dtype="float16")
    assert got.shape == (2, 5, 128)
    assert_close(got.reshape(10, 128), want, "float16", "bias")


def test_quantized_matmul_linearity_full_size(ops):
    """Size-independent property at the lm_head size (128256 x 4096): y(2x) == 2 y(x)
    exactly, and 16 sampled output columns match the oracle."""
    n, k = 128256, 4096
    g = torch.Generator(device="cuda").manual_seed(0)
    tw = torch.randint(-2**31, 2**31 - 1, (n, k // 8), generator=g, device="cuda", dtype=torch.int32)
    ts = (torch.rand((n, k // 64), generator=g, device="cuda") * 0.002 + 0.001).bfloat16()
    tb = (-8 * ts.float()).bfloat16()
    x = torch.randn((64, k), generator=g, device="cuda").bfloat16()
    y = ops.quantized_matmul(x, tw, ts, tb, group_size=64, bits=4)
    y2 = ops.quantized_matmul(x * 2, tw, ts, tb, group_size=64, bits=4)
    assert torch.equal(y2.float(), y.float() * 2)
    cols = torch.tensor([0, 1, 63, 64, 127, 128, 4095, 4096, 65535, 65536, 100000, 128000, 128128, 128254, 128255, 77],
                        device="cuda")
    want = oq.quantized_matmul(to_np(x), tw[cols].cpu().numpy().view(np.uint32), to_np(ts[cols]),
                               to_np(tb[cols]), 64, 4, dtype="bfloat16")
    assert_close(y[:, cols], want, "bfloat16", "sampled columns")


def test_quantized_matmul_validation(ops):
    _, _, _, tw, ts, tb = _quantized(64, 128, 4, 64)
    x = torch.zeros((2, 128), dtype=torch.bfloat16, device="cuda")
    with pytest.raises(ValueError):
        ops.quantized_matmul(x, tw, ts, tb, group_size=16, bits=4)
    with pytest.raises(ValueError):
        ops.quantized_matmul(x, tw, ts, tb, group_size=64, bits=3)
    with pytest.raises(ValueError):
        ops.quantized_matmul(x[:, :64], tw, ts, tb, group_size=64, bits=4)
    assert ops.quantized_matmul(x[:0], tw, ts, tb, group_size=64, bits=4).shape == (0, 64)


# ------------------------------------------------------- decoder neighbours ---

def test_rms_norm_and_residual(ops):
    x = torch.randn(7, 1024, device="cuda").bfloat16()
    w = torch.randn(1024, device="cuda").bfloat16()
    got = ops.rms_norm(x, w, 1e-6)
    xf = x.float()
    want = xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + 1e-6) * w.float()
    assert_close(got, to_np(want), "bfloat16", "rms_norm")
    res = torch.randn(7, 1024, device="cuda").bfloat16()
    r0 = res.clone()
    got = ops.rms_norm(x, w, 1e-6, residual=res)
    h = (x.float() + r0.float()).bfloat16()
    assert torch.equal(res, h)
    hf = h.float()
    want = hf * torch.rsqrt(hf.pow(2).mean(-1, keepdim=True) + 1e-6) * w.float()
    assert_close(got, to_np(want), "bfloat16", "rms_norm+residual")


@pytest.mark.parametrize("traditional", [False, True])
def test_rope(ops, traditional):
    t, hq, hk, d = 5, 8, 2, 128
    q = torch.randn(t, hq, d, device="cuda").bfloat16()
    k = torch.randn(t, hk, d, device="cuda").bfloat16()
    pos = torch.tensor([0, 1, 17, 4095, 8191], dtype=torch.int32, device="cuda")
    q0, k0 = q.clone(), k.clone()
    ops.rope(q, k, pos, d, 500000.0, traditional)
    inv = 500000.0 ** (-torch.arange(0, d, 2, device="cuda").float() / d)
    ang = pos.float()[:, None] * inv[None]
    cs, sn = ang.cos()[:, None], ang.sin()[:, None]

    def ref(x):
        x = x.float()
        if traditional:
            a, b = x[..., 0::2], x[..., 1::2]
            return torch.stack([a * cs - b * sn, b * cs + a * sn], -1).flatten(-2)
        a, b = x[..., : d // 2], x[..., d // 2:]
        return torch.cat([a * cs - b * sn, b * cs + a * sn], -1)

    assert_close(q, to_np(ref(q0)), "bfloat16", "rope q")
    assert_close(k, to_np(ref(k0)), "bfloat16", "rope k")


def test_silu_mul_argmax_add(ops):
    gu = torch.randn(6, 2 * 3072, device="cuda").bfloat16()
    got = ops.silu_mul(gu)
    g, u = gu[:, :3072].float(), gu[:, 3072:].float()
    want = torch.nn.functional.silu(g).bfloat16().float() * u
    assert_close(got, to_np(want), "bfloat16", "silu_mul")
    logits = torch.randn(9, 151936, device="cuda").bfloat16()
    logits[3, 777] = 100.0
    logits[3, 778] = 100.0  # tie -> first index, like mx.argmax
    got = ops.argmax(logits)
    assert got[3].item() == 777
    lf = logits.float()
    for r in range(9):
        i = got[r].item()
        assert lf[r, i] == lf[r].max() and (lf[r, :i] < lf[r, i]).all()  # first maximum
    y = torch.randn(4, 1000, device="cuda").bfloat16()
    x = torch.randn(4, 1000, device="cuda").bfloat16()
    y0 = y.clone()
    ops.add_(y, x)
    assert torch.equal(y, (y0.float() + x.float()).bfloat16())


@pytest.mark.parametrize("bits", [4, 8])
def test_prepack_round_trip_bit_exact(ops, bits):
    """b200_qlinear_prepack is a bijection on packed words (inverse restores the MLX layout)."""
    g = torch.Generator(device="cuda").manual_seed(bits)
    w = torch.randint(-2**31, 2**31 - 1, (257, 96), dtype=torch.int32, device="cuda", generator=g)
    wp = ops.qlinear_prepack(w, bits)
    assert wp.shape == w.shape
    assert torch.equal(ops.qlinear_prepack(wp, bits, inverse=True), w)
    if bits == 4:
        # nibbles (i, i+4) of the streaming word hold elements (2i, 2i+1) of the MLX word
        a, b = w.cpu().numpy().view(np.uint32), wp.cpu().numpy().view(np.uint32)
        for i in range(4):
            assert np.array_equal((b >> (4 * i)) & 0xF, (a >> (8 * i)) & 0xF)
            assert np.array_equal((b >> (4 * i + 16)) & 0xF, (a >> (8 * i + 4)) & 0xF)
    else:
        assert torch.equal(wp, w)


@pytest.mark.parametrize("variant", [0, 1, 2])
@pytest.mark.parametrize("m", [1, 16, 17, 64, 100, 300])
@pytest.mark.parametrize("n,k,bits,group", [(320, 1024, 4, 64), (256, 4096, 4, 64), (128, 2880, 4, 32),
                                           (1000, 2048, 4, 128), (384, 1536, 8, 64), (192, 4608, 8, 32)])
def test_prepacked_matmul_vs_oracle(ops, variant, m, n, k, bits, group):
    """Streaming layout + TMA-staged scales (variant 0/2) and per-thread scale loads (1): same
    results as the MLX-layout call (bit for bit) and within tolerance of the oracle."""
    ops.set_qlinear_variant(variant)
    try:
        wq, s, b, tw, ts, tb = _quantized(n, k, bits, group, seed=n + m)
        x = round_to(np.random.default_rng(m).standard_normal((m, k)).astype(np.float32), "bfloat16")
        tx = torch.from_numpy(x).bfloat16().cuda()
        bias = round_to(np.random.default_rng(7).standard_normal(n).astype(np.float32), "bfloat16")
        tbias = torch.from_numpy(bias).bfloat16().cuda()
        wp = ops.qlinear_prepack(tw, bits)
        got = ops.quantized_matmul(tx, wp, ts, tb, group_size=group, bits=bits, bias=tbias, prepacked=True)
        ops.set_qlinear_kernel(2)
        plain = ops.quantized_matmul(tx, tw, ts, tb, group_size=group, bits=bits, bias=tbias)
        want = oq.quantized_matmul(x, wq, s, b, group, bits, bias=bias, dtype="bfloat16")
        assert_close(got, want, "bfloat16", f"prepacked M={m} {n}x{k}")
        assert torch.equal(got, plain), "streaming layout changed the result"
    finally:
        ops.set_qlinear_kernel(0)
        ops.set_qlinear_variant(0)


def test_prepacked_rejects_unsupported(ops):
    """fp16 activations are not read by the tcgen05 kernel: loud error, no fallback."""
    wq, s, b, tw, ts, tb = _quantized(128, 512, 4, 64, "float16", seed=3)
    x = torch.randn((4, 512), device="cuda").half()
    with pytest.raises(ValueError, match="streaming layout"):
        ops.quantized_matmul(x, ops.qlinear_prepack(tw, 4), ts, tb, group_size=64, bits=4, prepacked=True)


def test_quantized_linear_module_prepack(ops):
    """prepack() picks the resident layout by size: small matrices keep the cluster split-K kernel's words
    (``weight_tc``), large ones the stream-K blob (``weight_blob``); either way ``mlx_weight()`` and checkpoints
    give the MLX layout back bit-exactly."""
    from vllm_metal_b200.quant.linear import QuantizedLinear

    for (n, k), attr in (((256, 512), "weight_tc"), ((8192 + 256, 4096), "weight_blob")):
        w = torch.randn((n, k), device="cuda").bfloat16() * 0.02
        lin = QuantizedLinear.from_dense(w, None, 64, 4, torch.bfloat16)
        x = torch.randn((8, k), device="cuda").bfloat16()
        y0 = lin(x)  # MLX-layout call
        lin.prepack()
        assert getattr(lin, attr, None) is not None and lin.weight.shape == (n, k // 8)
        assert (attr == "weight_blob") == (getattr(lin, "weight_tc", None) is None)
        y1 = lin(x)
        assert (y1.float() - y0.float()).abs().max().item() <= 2e-2 + 1e-2 * y0.float().abs().max().item()
        lin2 = QuantizedLinear.from_dense(w, None, 64, 4, torch.bfloat16).prepack(keep_mlx_layout=False)
        assert lin2.weight.shape[0] == 0 and torch.equal(lin2.mlx_weight(), lin.weight)
        assert torch.equal(lin2(x), y1)
        assert torch.equal(lin2.state_dict()["weight"], lin.weight)  # checkpoints keep the MLX layout
        # both layouts resident (prefill deployments): M > 64 takes the tensor-core GEMM, M <= 64 the blob
        lin3 = QuantizedLinear.from_dense(w, None, 64, 4, torch.bfloat16).prepack(keep_mlx_layout=False, prefill_layout=True)
        assert getattr(lin3, "weight_tc", None) is not None
        xb = torch.randn((200, k), device="cuda").bfloat16()
        yb = lin3(xb)
        assert (yb[:8].float() - lin3(xb[:8]).float()).abs().max().item() <= 2e-2 + 1e-2 * yb.float().abs().max().item()
