"""``QuantizedLinear`` / ``QuantizedEmbedding`` with the MLX module protocol.

The reference's hot-path linears are ``mlx.nn.QuantizedLinear`` objects whose
``__call__`` is ``mx.quantized_matmul(x, weight, scales, biases, transpose=True,
group_size, bits)`` (+ ``bias``); loaders and LoRA/AWQ code duck-type them by
``hasattr(bits) and hasattr(group_size)`` and read ``in = scales.shape[1] *
group_size``, ``out = scales.shape[0]`` (vllm_metal/quant/awq_loader.py:165-180,
v1/lora/layers.py:21-50).  These torch modules keep those attribute names and
shapes and call the sm_100a kernel through ``get_ops()``.
"""

from __future__ import annotations

import torch
from torch import nn

from ..metal import get_ops
from .affine import quantize


class QuantizedLinear(nn.Module):
    def __init__(self, in_features: int, out_features: int, bias: bool = False, group_size: int = 64,
                 bits: int = 4, dtype: torch.dtype = torch.bfloat16, device=None) -> None:
        super().__init__()
        if in_features % group_size:
            raise ValueError(f"in_features {in_features} not divisible by group_size {group_size}")
        self.group_size = group_size
        self.bits = bits
        words = in_features * bits // 32
        self.register_buffer("weight", torch.zeros((out_features, words), dtype=torch.int32, device=device))
        self.register_buffer("scales", torch.zeros((out_features, in_features // group_size), dtype=dtype, device=device))
        self.register_buffer("biases", torch.zeros((out_features, in_features // group_size), dtype=dtype, device=device))
        if bias:
            self.register_buffer("bias", torch.zeros((out_features,), dtype=dtype, device=device))
        else:
            self.bias = None
        self.swiglu = False

    @property
    def in_features(self) -> int:
        return self.scales.shape[1] * self.group_size

    @property
    def out_features(self) -> int:
        return self.scales.shape[0]

    @classmethod
    def from_dense(cls, weight: torch.Tensor, bias: torch.Tensor | None = None, group_size: int = 64,
                   bits: int = 4, dtype: torch.dtype = torch.bfloat16) -> QuantizedLinear:
        out_f, in_f = weight.shape
        m = cls(in_f, out_f, bias is not None, group_size, bits, dtype, weight.device)
        w, s, b = quantize(weight, group_size, bits, dtype)
        m.weight.copy_(w)
        m.scales.copy_(s)
        m.biases.copy_(b)
        if bias is not None:
            m.bias.copy_(bias.to(dtype))
        return m

    def shard_rows(self, rank: int, world: int) -> QuantizedLinear:
        """Column-parallel shard: rows (output features) [rank*N/world, ...)."""
        n = self.out_features
        if n % world:
            raise ValueError(f"out_features {n} not divisible by world size {world}")
        lo, hi = rank * n // world, (rank + 1) * n // world
        m = QuantizedLinear(self.in_features, hi - lo, self.bias is not None, self.group_size, self.bits,
                            self.scales.dtype, self.weight.device)
        m.weight.copy_(self.weight[lo:hi])
        m.scales.copy_(self.scales[lo:hi])
        m.biases.copy_(self.biases[lo:hi])
        if self.bias is not None:
            m.bias.copy_(self.bias[lo:hi])
        return m

    def shard_cols(self, rank: int, world: int) -> QuantizedLinear:
        """Row-parallel shard along K; K/world must be a multiple of group_size.
        The dense bias stays on rank 0 only (added once after the allreduce)."""
        k = self.in_features
        if k % world or (k // world) % self.group_size:
            raise ValueError(f"in_features {k} / {world} must be a multiple of group_size {self.group_size}")
        kw = k // world
        wl, wh = rank * kw * self.bits // 32, (rank + 1) * kw * self.bits // 32
        gl, gh = rank * kw // self.group_size, (rank + 1) * kw // self.group_size
        m = QuantizedLinear(kw, self.out_features, self.bias is not None and rank == 0, self.group_size,
                            self.bits, self.scales.dtype, self.weight.device)
        m.weight.copy_(self.weight[:, wl:wh])
        m.scales.copy_(self.scales[:, gl:gh])
        m.biases.copy_(self.biases[:, gl:gh])
        if m.bias is not None:
            m.bias.copy_(self.bias)
        return m

    #: rows of x above which the tensor-core GEMM layout (``weight_tc``) is used when both layouts are resident
    PREFILL_M = 64
    #: packed-weight bytes up to which a (non-SwiGLU) linear is served by the cluster split-K kernel alone.
    #: Measured at M = 64 (profiles/r02_qstream_vs_cluster_tp_shapes.txt): below ~16 MB a launch is all fixed
    #: cost, and the cluster kernel's DSMEM reduction (5-11 us) beats the stream-K kernel's exchange through
    #: global memory (8-13 us); from ~18 MB up the persistent stream wins (and it fuses SwiGLU).
    SMALL_BYTES = 16 << 20

    def prepack(self, keep_mlx_layout: bool = True, swiglu: bool = False, prefill_layout: bool = False) -> QuantizedLinear:
        """Load-time transform into the kernels' layouts; picks per shape:

        * ``weight_blob`` -- streaming blob of the persistent stream-K kernel (``b200_qstream_prepack``).
          ``swiglu=True``: rows are ``[gate; up]`` and the call returns ``silu(gate) * up``
          (``[.., out_features / 2]``), the fusion of the reference's compiled_mlp.py.
        * ``weight_tc`` -- the words the tcgen05 cluster split-K / M > 64 GEMM kernel (csrc/qgemm_tc.cu) reads
          (nibble-permuted for int4, the MLX words for int8).  Kept INSTEAD of the blob for small matrices
          (``SMALL_BYTES``), and BESIDE it with ``prefill_layout=True``: the stream-K kernel re-streams and
          re-dequantises the weights once per 64 rows of x, which a 2048-token prefill chunk cannot afford.

        ``weight`` keeps the MLX layout for loaders / LoRA introspection unless ``keep_mlx_layout=False``
        (then ``mlx_weight()`` rebuilds it bit-exactly from whichever layout is resident)."""
        if self.scales.dtype not in (torch.bfloat16, torch.float16) or not self.weight.shape[0]:
            return self
        ops = get_ops()
        tc_ok = self.scales.dtype == torch.bfloat16 and self.in_features % 64 == 0
        small = tc_ok and not swiglu and self.weight.numel() * 4 <= self.SMALL_BYTES
        blob = None
        if not small:
            try:
                blob = ops.qstream_prepack(self.weight, self.scales, self.biases, self.group_size, self.bits, swiglu)
            except ValueError:
                blob = None  # a scale/bias outside fp16: stay on the MLX-layout kernels
        if tc_ok and (small or prefill_layout or blob is None):
            tc = ops.qlinear_prepack(self.weight, self.bits) if self.bits == 4 else self.weight.clone()
            self.register_buffer("weight_tc", tc, persistent=False)
        if blob is not None:
            self.register_buffer("weight_blob", blob, persistent=False)
            self.swiglu = bool(swiglu)
        if not keep_mlx_layout and (blob is not None or getattr(self, "weight_tc", None) is not None):
            self.weight = torch.empty((0, self.weight.shape[1]), dtype=torch.int32, device=self.scales.device)
        return self

    def mlx_weight(self) -> torch.Tensor:
        if self.weight.shape[0]:
            return self.weight
        if getattr(self, "weight_blob", None) is not None:
            return get_ops().qstream_unpack(self.weight_blob, self.out_features, self.in_features, self.group_size,
                                            self.bits, self.swiglu)
        tc = getattr(self, "weight_tc", None)
        if tc is not None:
            return get_ops().qlinear_prepack(tc, self.bits, inverse=True) if self.bits == 4 else tc
        return self.weight

    def _save_to_state_dict(self, destination, prefix, keep_vars):
        # checkpoints always carry the MLX layout, whatever the resident layout is
        super()._save_to_state_dict(destination, prefix, keep_vars)
        if prefix + "weight" in destination and destination[prefix + "weight"].shape[0] == 0:
            destination[prefix + "weight"] = self.mlx_weight()

    def forward(self, x: torch.Tensor, out_f16: bool = False) -> torch.Tensor:
        """``x`` in the module dtype, or fp16 (the stream-K kernel's operand type: a bf16 ``x`` is staged as fp16
        by one extra launch, so producers that can emit fp16 do).  ``out_f16``: a bf16 module stores its
        (bf16-rounded) result as fp16 for a following QuantizedLinear."""
        blob = getattr(self, "weight_blob", None)
        tc = getattr(self, "weight_tc", None)
        if tc is not None and x.dtype in (torch.bfloat16, torch.float16) and (
                blob is None or x.numel() // x.shape[-1] > self.PREFILL_M):
            y = get_ops().quantized_matmul(x, tc, self.scales, self.biases, transpose=True, group_size=self.group_size,
                                           bits=self.bits, bias=self.bias, prepacked=True)
            if self.swiglu:  # the blob's fused epilogue, unfused here
                y = get_ops().silu_mul(y)
            return y
        if blob is not None and x.dtype in (torch.bfloat16, torch.float16):
            dt = self.scales.dtype
            return get_ops().qstream_linear(x, blob, self.out_features, self.in_features, self.group_size, self.bits,
                                            bias=self.bias, swiglu=self.swiglu,
                                            out_dtype=torch.float16 if out_f16 else dt, round_dtype=dt)
        if x.dtype != self.scales.dtype:
            x = x.to(self.scales.dtype)
        return get_ops().quantized_matmul(x, self.mlx_weight(), self.scales, self.biases, transpose=True,
                                          group_size=self.group_size, bits=self.bits, bias=self.bias)


class QuantizedEmbedding(nn.Module):
    """Quantized token embedding; ``as_linear`` is the tied lm_head
    (reference: gguf/mlx_native.py:224-240, gguf/wrappers.py:88-116)."""

    def __init__(self, num_embeddings: int, dims: int, group_size: int = 64, bits: int = 4,
                 dtype: torch.dtype = torch.bfloat16, device=None) -> None:
        super().__init__()
        self.group_size = group_size
        self.bits = bits
        self.register_buffer("weight", torch.zeros((num_embeddings, dims * bits // 32), dtype=torch.int32, device=device))
        self.register_buffer("scales", torch.zeros((num_embeddings, dims // group_size), dtype=dtype, device=device))
        self.register_buffer("biases", torch.zeros((num_embeddings, dims // group_size), dtype=dtype, device=device))

    @classmethod
    def from_dense(cls, weight: torch.Tensor, group_size: int = 64, bits: int = 4,
                   dtype: torch.dtype = torch.bfloat16) -> QuantizedEmbedding:
        n, d = weight.shape
        m = cls(n, d, group_size, bits, dtype, weight.device)
        w, s, b = quantize(weight, group_size, bits, dtype)
        m.weight.copy_(w)
        m.scales.copy_(s)
        m.biases.copy_(b)
        return m

    def forward(self, ids: torch.Tensor) -> torch.Tensor:
        return get_ops().dequantize(self.weight, self.scales, self.biases, self.group_size, self.bits,
                                    ids=ids, dtype=self.scales.dtype)

    head_rows: tuple[int, int] | None = None  # vocab-parallel tied head: the rows ``as_linear`` multiplies by

    def prepack(self, rows: tuple[int, int] | None = None) -> QuantizedEmbedding:
        """Streaming blob for the tied lm_head (``as_linear``); the row gather keeps the MLX layout.  ``rows`` =
        (lo, hi): a vocab-parallel rank's slice of the head -- ``as_linear`` then returns logits [.., hi - lo]."""
        self.head_rows = rows
        lo, hi = rows if rows is not None else (0, self.scales.shape[0])
        if self.scales.dtype in (torch.bfloat16, torch.float16):
            try:
                blob = get_ops().qstream_prepack(self.weight[lo:hi], self.scales[lo:hi], self.biases[lo:hi],
                                                 self.group_size, self.bits)
            except ValueError:
                return self
            self.register_buffer("weight_blob", blob, persistent=False)
        return self

    def as_linear(self, x: torch.Tensor) -> torch.Tensor:
        blob = getattr(self, "weight_blob", None)
        lo, hi = self.head_rows if self.head_rows is not None else (0, self.scales.shape[0])
        k = self.scales.shape[1] * self.group_size
        if blob is not None and x.dtype in (torch.bfloat16, torch.float16):
            return get_ops().qstream_linear(x, blob, hi - lo, k, self.group_size, self.bits, out_dtype=self.scales.dtype)
        if x.dtype != self.scales.dtype:
            x = x.to(self.scales.dtype)
        return get_ops().quantized_matmul(x, self.weight[lo:hi], self.scales[lo:hi], self.biases[lo:hi], transpose=True,
                                          group_size=self.group_size, bits=self.bits)
