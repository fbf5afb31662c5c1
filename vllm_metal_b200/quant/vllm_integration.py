"""Quant-method hook: marker quantization config for MLX-style affine group quantisation.

Mirrors the mechanism of ``vllm_metal/gguf/vllm_integration.py:46-118``: vLLM bare-constructs the
registered ``QuantizationConfig`` and calls its classmethods on every serve, but the plugin's own
runner loads and runs the quantized linears (``quant/linear.py::QuantizedLinear`` ->
``b200_qlinear``), so the config only has to validate ``quantization="mlx_affine"`` and carry
``bits`` / ``group_size``.  Mounted through the ``vllm.general_plugins`` entry point
(``vllm_metal_b200:register_quant_config``).
"""

from __future__ import annotations

import torch
from vllm.model_executor.layers.quantization import (
    QUANTIZATION_METHODS,
    get_quantization_config,
    register_quantization_config,
)
from vllm.model_executor.layers.quantization.base_config import QuantizationConfig

NAME = "mlx_affine"


class MLXAffineConfig(QuantizationConfig):
    """``{"quantization": {"group_size": 32|64|128, "bits": 4|8}}`` as mlx-lm writes it into
    ``config.json``.  Lives at module top level: it is pickled into the
    EngineCore / worker processes."""

    _SUPPORTED_ACT_DTYPES = [torch.bfloat16, torch.float16]

    def __init__(self, bits: int = 4, group_size: int = 64) -> None:
        super().__init__()
        if bits not in (4, 8):
            raise ValueError(f"mlx_affine: bits must be 4 or 8, got {bits}")
        if group_size not in (32, 64, 128):
            raise ValueError(f"mlx_affine: group_size must be 32, 64 or 128, got {group_size}")
        self.bits, self.group_size = bits, group_size

    @classmethod
    def get_name(cls) -> str:
        return NAME

    @classmethod
    def get_supported_act_dtypes(cls) -> list[torch.dtype]:
        return cls._SUPPORTED_ACT_DTYPES

    @classmethod
    def get_min_capability(cls) -> int:
        return 100  # sm_100a kernels only

    @classmethod
    def get_config_filenames(cls) -> list[str]:
        return []

    @classmethod
    def from_config(cls, config: dict) -> "MLXAffineConfig":
        q = config.get("quantization", config) if isinstance(config, dict) else {}
        return cls(bits=int(q.get("bits", 4)), group_size=int(q.get("group_size", 64)))

    def get_quant_method(self, layer: torch.nn.Module, prefix: str):
        # never exercised: the plugin bypasses vLLM's model executor (reference behaviour)
        return None


def register_mlx_affine_config() -> None:
    """Idempotent registration (reference ``GGUFEngineIntegration.register``, :99-118)."""
    if NAME not in QUANTIZATION_METHODS or get_quantization_config(NAME) is not MLXAffineConfig:
        register_quantization_config(NAME)(MLXAffineConfig)
