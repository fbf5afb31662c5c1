"""MLX-affine quantized linears on B200 (reference: vllm_metal/quant + MLX core)."""
from .linear import QuantizedEmbedding, QuantizedLinear  # noqa: F401
