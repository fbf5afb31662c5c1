"""B200-native decode hot path of vllm-metal, behind the reference's plugin surface.

Entry points mirror ``vllm_metal/__init__.py:117-145``: ``register()`` is the
``vllm.platform_plugins`` hook and returns the dotted path of the ``Platform``
subclass when a CUDA device of compute capability 10.x is present (``None``
otherwise, so vLLM falls through to its other platforms).  Importing this
package does not import vLLM or load the CUDA library; ``metal.get_ops()`` does
the latter on first use and raises if it is missing (no CPU fallback).
"""

from __future__ import annotations

__version__ = "0.1.0"

PLATFORM_CLS = "vllm_metal_b200.platform.B200Platform"


def _cuda_sm100_available() -> bool:
    try:
        import torch

        if not torch.cuda.is_available():
            return False
        major, _ = torch.cuda.get_device_capability(0)
        return major == 10
    except Exception:  # noqa: BLE001
        return False


def register() -> str | None:
    """``vllm.platform_plugins`` entry point (reference ``_register``, __init__.py:117-145).

    Claims the platform only on an sm_100 device AND when ``VLLM_B200_PLUGIN=1`` is set: vLLM's in-tree
    CUDA platform also runs on a B200, so taking the device over must be the operator's explicit choice."""
    import os

    if os.environ.get("VLLM_B200_PLUGIN", "0") in ("", "0"):
        return None
    return PLATFORM_CLS if _cuda_sm100_available() else None


def register_quant_config() -> None:
    """``vllm.general_plugins`` entry point: registers the marker quantization
    config (reference gguf/vllm_integration.py:110-118)."""
    from .quant.vllm_integration import register_mlx_affine_config

    register_mlx_affine_config()
