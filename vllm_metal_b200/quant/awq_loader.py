"""AWQ load path: AutoAWQ GEMM triples -> MLX-affine ``QuantizedLinear`` modules.

Reference: ``vllm_metal/quant/awq_loader.py:49-224`` (``AWQQuantLoader``: preflight via
``normalize_quant_config``, GPTQ refused, the repack delegated to
``mlx_lm.utils._transform_awq_weights`` -- third-party, mlx-lm @ 254d153 --, then every
non-quantized floating parameter aligned to the runtime dtype while the transform's
``scales`` / ``biases`` keep theirs).  mlx-lm is not available here, so the repack is
restated from the contract the reference pins (``tests/quant/test_awq_repack_parity.py``):

* ``qweight`` int32 ``[in, out/8]``, ``qzeros`` int32 ``[groups, out/8]`` in AutoAWQ's
  interleaved nibble order ``(0,2,4,6,1,3,5,7)``, ``scales`` fp16 ``[groups, out]``;
* result ``weight`` uint32 ``[out, in/8]`` (MLX order: element k at bits ``4*(k%8)``),
  ``scales`` ``[out, groups]``, ``biases = -(zeros * scales)`` ``[out, groups]``, so that
  ``scales*q + biases == (q - z)*s``.

The repack is integer work on whatever device the tensors live on (torch, load time);
the decode step never sees AWQ's layout.
"""

from __future__ import annotations

import json
from collections.abc import Mapping
from pathlib import Path
from typing import Any

import torch
from torch import nn

from .awq_config import UnsupportedQuantizationConfigError, normalize_quant_config
from .linear import QuantizedLinear

AWQ_PACK_ORDER = (0, 2, 4, 6, 1, 3, 5, 7)  # nibble j of a word holds column AWQ_PACK_ORDER[j] of its 8
# Buffers whose dtype belongs to the transform, not to the runtime (awq_loader.py:40-43).
_QUANT_BUFFER_NAMES = ("scales", "biases")


def _awq_unpack(packed: torch.Tensor) -> torch.Tensor:
    """int32 ``[..., C]`` in AutoAWQ order -> uint8 nibbles ``[..., 8*C]`` in natural column order."""
    u = packed.to(torch.int64) & 0xFFFFFFFF
    out = torch.empty(packed.shape[:-1] + (packed.shape[-1] * 8,), dtype=torch.uint8, device=packed.device)
    for shift_idx, pos in enumerate(AWQ_PACK_ORDER):
        out[..., pos::8] = ((u >> (4 * shift_idx)) & 0xF).to(torch.uint8)
    return out


def _mlx_pack4(q: torch.Tensor) -> torch.Tensor:
    """uint8 nibbles ``[N, K]`` -> MLX int4 words ``[N, K/8]`` (int32 bit patterns)."""
    n, k = q.shape
    u = q.to(torch.int64).reshape(n, k // 8, 8)
    shifts = torch.arange(8, device=q.device, dtype=torch.int64) * 4
    word = (u << shifts).sum(dim=-1)
    return torch.where(word >= 2**31, word - 2**32, word).to(torch.int32)


def transform_awq_weights(weights: Mapping[str, torch.Tensor], quantization_config: Mapping[str, Any]):
    """``{"<p>.qweight", "<p>.qzeros", "<p>.scales"}`` -> ``{"<p>.weight", "<p>.scales", "<p>.biases"}``.

    Other entries pass through.  Returns ``(new_weights, mlx_quant_config)`` like
    ``mlx_lm.utils._transform_awq_weights``.  ``*.g_idx`` (GPTQ act-order) is refused."""
    cfg = normalize_quant_config(quantization_config)
    out: dict[str, torch.Tensor] = {}
    for name in weights:
        if name.endswith(".g_idx"):
            raise ValueError(f"{name}: g_idx (activation-ordered GPTQ) checkpoints are not supported")
    for name, t in weights.items():
        if name.endswith(".qweight"):
            prefix = name[: -len(".qweight")]
            qzeros, scales = weights[prefix + ".qzeros"], weights[prefix + ".scales"]
            q = _awq_unpack(t)  # [in, out]
            z = _awq_unpack(qzeros)  # [groups, out]
            in_f, out_f = q.shape
            if in_f % cfg["group_size"] or scales.shape != (in_f // cfg["group_size"], out_f) or z.shape != scales.shape:
                raise ValueError(f"{prefix}: inconsistent AWQ shapes qweight {tuple(t.shape)}, "
                                 f"qzeros {tuple(qzeros.shape)}, scales {tuple(scales.shape)}")
            out[prefix + ".weight"] = _mlx_pack4(q.t().contiguous())
            s_t = scales.t().contiguous()
            out[prefix + ".scales"] = s_t
            # -(z*s) in the scales' own dtype: the one extra rounding the reference's test allows for
            out[prefix + ".biases"] = -(z.t().contiguous().to(s_t.dtype) * s_t)
        elif name.endswith((".qzeros", ".scales")) and (name.rsplit(".", 1)[0] + ".qweight") in weights:
            continue
        else:
            out[name] = t
    return out, {"group_size": cfg["group_size"], "bits": cfg["bits"]}


class AWQQuantLoader:
    """Owner of the AWQ branch of model loading (reference ``AWQQuantLoader``)."""

    def __init__(self, normalized_quant_config: Mapping[str, Any]) -> None:
        self.quantization_config = dict(normalized_quant_config)

    @classmethod
    def for_model(cls, model_name: str) -> AWQQuantLoader | None:
        """Inspect ``<model_name>/config.json``; ``None`` when the checkpoint is not AWQ.
        GPTQ is refused outright (awq_loader.py:84-92); only local directories are read
        (there is no network here -- a Hub id raises instead of silently skipping the preflight)."""
        raw = cls._read_raw_quantization_config(model_name)
        if raw is None:
            return None
        method = raw.get("quant_method")
        if method == "gptq":
            raise UnsupportedQuantizationConfigError(
                "GPTQ checkpoints are not validated for this build; only AWQ is. Use an AWQ release of "
                "this checkpoint, or re-quantize with AWQ.")
        if method != "awq":
            return None
        return cls(normalize_quant_config(raw))

    @staticmethod
    def _read_raw_quantization_config(model_name: str) -> Mapping[str, Any] | None:
        path = Path(model_name)
        if not path.is_dir():
            raise FileNotFoundError(f"{model_name}: not a local checkpoint directory (no Hub access in this build)")
        cfg_path = path / "config.json"
        if not cfg_path.is_file():
            return None
        config = json.loads(cfg_path.read_text(encoding="utf-8"))
        qc = config.get("quantization_config")
        if isinstance(qc, dict):
            return qc
        nested = config.get("text_config")
        if isinstance(nested, dict) and isinstance(nested.get("quantization_config"), dict):
            return nested["quantization_config"]
        return None

    def linear_from_awq(self, qweight, qzeros, scales, bias=None) -> QuantizedLinear:
        """One AWQ layer -> a ``QuantizedLinear`` holding the MLX-affine triple (scales/biases keep
        the checkpoint's dtype, fp16 for AutoAWQ)."""
        w, _ = transform_awq_weights({"l.qweight": qweight, "l.qzeros": qzeros, "l.scales": scales},
                                     self.quantization_config)
        out_f, words = w["l.weight"].shape
        m = QuantizedLinear(words * 8, out_f, bias is not None, self.quantization_config["group_size"],
                            self.quantization_config["bits"], w["l.scales"].dtype, qweight.device)
        m.weight.copy_(w["l.weight"])
        m.scales.copy_(w["l.scales"])
        m.biases.copy_(w["l.biases"])
        if bias is not None:
            m.bias.copy_(bias.to(m.bias.dtype))
        return m

    @staticmethod
    def _is_quantized_module(module: Any) -> bool:
        # the MLX quantize protocol is duck-typed (awq_loader.py:165-180)
        return hasattr(module, "bits") and hasattr(module, "group_size")

    @classmethod
    def align_non_quantized_dtypes(cls, model: nn.Module, target_dtype: torch.dtype) -> int:
        """Cast floating tensors of leaf modules to ``target_dtype`` except the transform-owned
        ``scales`` / ``biases`` of quantized leaves; returns how many were cast
        (reference ``_align_non_quantized_dtypes``, awq_loader.py:182-224)."""
        n_cast = 0
        for module in model.modules():
            if any(True for _ in module.children()):
                continue
            quantized = cls._is_quantized_module(module)
            named = list(module.named_parameters(recurse=False)) + list(module.named_buffers(recurse=False))
            for name, value in named:
                if quantized and name in _QUANT_BUFFER_NAMES:
                    continue
                if value is None or not value.dtype.is_floating_point or value.dtype == target_dtype:
                    continue
                value.data = value.data.to(target_dtype)
                n_cast += 1
        return n_cast
