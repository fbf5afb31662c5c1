"""The reference's OWN pins for the quantized-linear path, run through the CUDA path.

MLX (where ``mx.quantize/dequantize/quantized_matmul`` live) is third-party and absent, so
the only reference-held anchors for rows a13-a15 are these two test files; both are ported
here with their tolerances unchanged and their synthetic inputs regenerated from the same
seeds:

* ``/root/reference/tests/quant/test_awq_repack_parity.py:100-176`` -- an AutoAWQ triple
  (seed 42; ``W = (q - z) * s``) goes through the product's AWQ transform and
  ``b200_dequantize_rows``; max error <= ``max(max|W| * 2**-9, 2e-3)``, mean < a quarter of it,
  shapes ``weight [out, in/8]`` uint32, ``scales/biases [out, in/128]``.
* ``/root/reference/tests/test_gguf_mlx_native.py:70-86`` -- group-32 ``matmul`` with float32
  activations against the dense dequantised oracle at ``atol = max|ref| * 4e-6``; output dtype
  follows ``x`` (:89-99); leading dims preserved (:102-110).
"""

from __future__ import annotations

import numpy as np
import pytest
import torch

from oracle import affine_quant as oq

from .gpu_helpers import to_np

pytestmark = pytest.mark.gpu

GROUP_SIZE, BITS = 128, 4


def _build_awq_layer(in_features, out_features, rng):
    n_groups = in_features // GROUP_SIZE
    q = rng.integers(0, 16, size=(in_features, out_features), dtype=np.uint8)
    z = rng.integers(0, 16, size=(n_groups, out_features), dtype=np.uint8)
    s = rng.uniform(0.001, 0.05, size=(n_groups, out_features)).astype(np.float16)
    ref = np.empty((in_features, out_features), np.float32)
    for g in range(n_groups):
        sl = slice(g * GROUP_SIZE, (g + 1) * GROUP_SIZE)
        ref[sl] = (q[sl].astype(np.float32) - z[g].astype(np.float32)) * s[g].astype(np.float32)
    return oq.awq_pack_int32(q), oq.awq_pack_int32(z), s, ref.T


@pytest.mark.parametrize(("in_features", "out_features"), [(128, 64), (128, 256), (256, 256)],
                         ids=["one-group-narrow", "one-group-wide", "two-groups-square"])
def test_awq_repack_dequant_matches_awq_ground_truth(ops, in_features, out_features):
    from vllm_metal_b200.quant.awq_loader import transform_awq_weights

    rng = np.random.default_rng(seed=42)
    qweight, qzeros, scales, ref_w = _build_awq_layer(in_features, out_features, rng)
    weights = {"blk.0.qweight": torch.from_numpy(qweight).cuda(), "blk.0.qzeros": torch.from_numpy(qzeros).cuda(),
               "blk.0.scales": torch.from_numpy(scales).cuda()}
    new, cfg = transform_awq_weights(
        weights, {"quant_method": "awq", "bits": BITS, "group_size": GROUP_SIZE, "zero_point": True, "version": "gemm"})
    assert {"blk.0.weight", "blk.0.scales", "blk.0.biases"} <= set(new)
    assert cfg["bits"] == BITS and cfg["group_size"] == GROUP_SIZE
    n_groups = in_features // GROUP_SIZE
    assert new["blk.0.weight"].shape == (out_features, in_features // 8)
    assert new["blk.0.scales"].shape == (out_features, n_groups)
    assert new["blk.0.biases"].shape == (out_features, n_groups)
    # the oracle's restatement of the same repack agrees bit for bit on the integer part
    ow, os_, _ = oq.awq_to_mlx_affine(qweight, qzeros, scales)
    assert np.array_equal(new["blk.0.weight"].cpu().numpy().view(np.uint32), ow)
    assert np.array_equal(to_np(new["blk.0.scales"]), os_)

    deq = to_np(ops.dequantize(new["blk.0.weight"], new["blk.0.scales"], new["blk.0.biases"], GROUP_SIZE, BITS,
                               dtype=torch.float32))
    diff = np.abs(deq - ref_w)
    ceiling = max(np.abs(ref_w).max() * 2.0 ** -9, 2e-3)
    assert diff.max() <= ceiling, f"max dequant err {diff.max():.4e} > {ceiling:.4e}"
    assert diff.mean() < ceiling / 4

    # and the matmul over the same triple (fp16 activations, AWQ's fp16 scales) vs x @ W^T
    x = np.random.default_rng(1).standard_normal((5, in_features)).astype(np.float16)
    y = ops.quantized_matmul(torch.from_numpy(x).cuda(), new["blk.0.weight"], new["blk.0.scales"], new["blk.0.biases"],
                             group_size=GROUP_SIZE, bits=BITS)
    want = x.astype(np.float32) @ ref_w.T
    assert y.dtype == torch.float16
    assert np.abs(to_np(y) - want).max() <= 1.5e-2 + 2e-2 * np.abs(want).max()


def test_awq_transform_rejects_g_idx_and_bad_configs():
    from vllm_metal_b200.quant.awq_config import UnsupportedQuantizationConfigError
    from vllm_metal_b200.quant.awq_loader import transform_awq_weights

    w = {"b.qweight": torch.zeros((128, 8), dtype=torch.int32), "b.qzeros": torch.zeros((1, 8), dtype=torch.int32),
         "b.scales": torch.zeros((1, 64), dtype=torch.float16), "b.g_idx": torch.zeros(128, dtype=torch.int32)}
    with pytest.raises(ValueError, match="g_idx"):
        transform_awq_weights(w, {"quant_method": "awq", "bits": 4, "group_size": 128})
    with pytest.raises(UnsupportedQuantizationConfigError):
        transform_awq_weights({}, {"quant_method": "gptq", "bits": 4, "group_size": 128})


def _g32_tensor(bits, shape=(64, 128)):
    w = np.random.default_rng(0).standard_normal(shape).astype(np.float32)
    wq, s, b = oq.quantize(w, 32, bits, "float16")  # GGUF-native tensors are g=32, f16 scales
    dense = oq.dequantize(wq, s, b, 32, bits, "float32")
    return (torch.from_numpy(wq.view(np.int32)).cuda(), torch.from_numpy(s).half().cuda(),
            torch.from_numpy(b).half().cuda(), dense)


@pytest.mark.parametrize("bits", [4, 8])
def test_matmul_matches_dense_oracle_f32(ops, bits):
    tw, ts, tb, dense = _g32_tensor(bits)
    x = np.random.default_rng(1).standard_normal((3, 128)).astype(np.float32)
    out = ops.quantized_matmul(torch.from_numpy(x).cuda(), tw, ts, tb, group_size=32, bits=bits)
    expected = x.astype(np.float64) @ dense.astype(np.float64).T
    assert out.dtype == torch.float32
    np.testing.assert_allclose(to_np(out), expected, rtol=0, atol=float(np.abs(expected).max()) * 4e-6)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16, torch.bfloat16])
@pytest.mark.parametrize("bits", [4, 8])
def test_matmul_output_dtype_follows_x_and_keeps_leading_shape(ops, bits, dtype):
    tw, ts, tb, dense = _g32_tensor(bits)
    x = torch.randn((2, 3, 128), device="cuda").to(dtype)
    out = ops.quantized_matmul(x, tw, ts, tb, group_size=32, bits=bits)
    assert out.dtype == dtype and out.shape == (2, 3, 64)
    want = to_np(x).reshape(6, 128) @ dense.T
    tol = 1e-4 if dtype == torch.float32 else 3e-2
    assert np.abs(to_np(out).reshape(6, 64) - want).max() <= tol + tol * np.abs(want).max()
