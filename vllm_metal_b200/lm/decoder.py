"""Decoder-only transformer over the paged ops (Llama / Qwen3 / gpt-oss attention).

In the reference the layer wiring is third-party ``mlx_lm`` model code whose
``self_attn`` modules are replaced by the paged-attention wrapper
(vllm_metal/attention/impls/sdpa_wrapper.py:45-191); this file is the torch
stand-in with the same attribute contract (``q_proj/k_proj/v_proj/o_proj`` or
packed ``qkv_proj``, ``n_heads``, ``n_kv_heads``, ``scale``, ``q_norm``/
``k_norm``, ``rope``, ``sinks``) so ``sdpa_forward`` duck-types it identically.
Weights are MLX-affine quantized (``QuantizedLinear``); tensor parallelism shards
heads / MLP columns per rank and sums o_proj / down_proj outputs with an
allreduce.
"""

from __future__ import annotations

from dataclasses import dataclass, field

import torch
from torch import nn

from ..attention.impls.sdpa import sdpa_forward
from ..metal import get_ops
from ..quant.linear import QuantizedEmbedding, QuantizedLinear


@dataclass
class ModelArgs:
    name: str
    num_layers: int
    hidden: int
    n_heads: int
    n_kv_heads: int
    head_dim: int
    intermediate: int
    vocab: int
    rms_eps: float = 1e-5
    rope_base: float = 500000.0
    qk_norm: bool = False
    tie_embeddings: bool = False
    sinks: bool = False
    sliding_window: int = -1
    # "full" | "sliding_attention" per layer (reference: build_sliding_window_per_layer,
    # vllm_metal/v1/model_adapter.py:693-710)
    layer_types: list[str] = field(default_factory=list)
    bits: int = 4
    group_size: int = 64
    mlp: bool = True  # False: attention block only (gpt-oss config: MoE MLP out of scope)
    # mixture-of-experts MLP (quant/switch_linear.py): > 0 experts replaces the dense MLP of every layer
    num_experts: int = 0
    experts_per_tok: int = 0
    moe_activation: str = "silu"        # "silu" | "gpt_oss" (clipped SwiGLU)
    moe_routing: str = "topk_softmax"   # gpt-oss; "softmax_topk" = Qwen3-MoE
    moe_bias: bool = False

    def sliding_window_per_layer(self) -> list[int]:
        if not self.layer_types:
            return [-1] * self.num_layers
        return [self.sliding_window if t == "sliding_attention" else -1 for t in self.layer_types]


def _gpt_oss_layers(n: int) -> list[str]:
    return ["sliding_attention" if i % 2 == 0 else "full_attention" for i in range(n)]


PRESETS = {
    "qwen3-0.6b": ModelArgs("Qwen3-0.6B", 28, 1024, 16, 8, 128, 3072, 151936, 1e-6, 1e6, True, True),
    "llama-3.1-8b": ModelArgs("Llama-3.1-8B", 32, 4096, 32, 8, 128, 14336, 128256),
    "qwen3-32b": ModelArgs("Qwen3-32B", 64, 5120, 64, 8, 128, 25600, 151936, 1e-6, 1e6, True, False, bits=8),
    "llama-3.1-70b": ModelArgs("Llama-3.1-70B", 80, 8192, 64, 8, 128, 28672, 128256),
    "gpt-oss-20b-attn": ModelArgs("gpt-oss-20b (attention blocks)", 24, 2880, 64, 8, 64, 2880, 201088, 1e-5,
                                  150000.0, False, False, True, 128, _gpt_oss_layers(24), mlp=False),
    "tiny": ModelArgs("tiny-test", 2, 256, 8, 2, 64, 512, 1024, 1e-6, 10000.0, True, True),
}


class RoPE:
    """Parameters of ``mlx.nn.RoPE`` (dims, base, traditional); applied by ``ops.rope``."""

    def __init__(self, dims: int, base: float, traditional: bool = False) -> None:
        self.dims, self.base, self.traditional = dims, base, traditional


class RMSNorm(nn.Module):
    def __init__(self, dims: int, eps: float, dtype, device) -> None:
        super().__init__()
        self.register_buffer("weight", torch.ones(dims, dtype=dtype, device=device))
        self.eps = eps

    def forward(self, x: torch.Tensor, residual: torch.Tensor | None = None, out_f16: bool = False) -> torch.Tensor:
        """``out_f16``: the (bf16-rounded) result stored as fp16, the operand type of the stream-K linears."""
        f16 = out_f16 and x.dtype == torch.bfloat16 and x.shape[-1] % 8 == 0 and 512 <= x.shape[-1] <= 16384
        return get_ops().rms_norm(x, self.weight, self.eps, residual=residual, out_dtype=torch.float16 if f16 else None)


def _rand_qlinear(in_f, out_f, args: ModelArgs, gen, dtype, device, std=0.02) -> QuantizedLinear:
    w = torch.empty((out_f, in_f), dtype=torch.float32, device=device).normal_(0.0, std, generator=gen)
    return QuantizedLinear.from_dense(w.to(dtype), None, args.group_size, args.bits, dtype)


def _cat_rows(mods: list[QuantizedLinear]) -> QuantizedLinear:
    """Fuse linears that share an input by stacking their output rows (row-wise
    quantisation makes this exact)."""
    m0 = mods[0]
    out = QuantizedLinear(m0.in_features, sum(m.out_features for m in mods), False, m0.group_size, m0.bits,
                          m0.scales.dtype, m0.weight.device)
    out.weight.copy_(torch.cat([m.weight for m in mods], 0))
    out.scales.copy_(torch.cat([m.scales for m in mods], 0))
    out.biases.copy_(torch.cat([m.biases for m in mods], 0))
    return out


class Attention(nn.Module):
    def __init__(self, args: ModelArgs, gen, dtype, device, rank: int, world: int) -> None:
        super().__init__()
        if args.n_kv_heads % world:
            raise ValueError(f"n_kv_heads {args.n_kv_heads} not divisible by TP size {world}")
        self.n_heads = args.n_heads // world
        self.n_kv_heads = args.n_kv_heads // world
        self.head_dim = args.head_dim
        self.scale = args.head_dim ** -0.5
        d = args.head_dim
        q = _rand_qlinear(args.hidden, args.n_heads * d, args, gen, dtype, device).shard_rows(rank, world)
        k = _rand_qlinear(args.hidden, args.n_kv_heads * d, args, gen, dtype, device).shard_rows(rank, world)
        v = _rand_qlinear(args.hidden, args.n_kv_heads * d, args, gen, dtype, device).shard_rows(rank, world)
        self.qkv_proj = _cat_rows([q, k, v])
        self.o_proj = _rand_qlinear(args.n_heads * d, args.hidden, args, gen, dtype, device).shard_cols(rank, world)
        if args.qk_norm:
            self.q_norm = RMSNorm(d, args.rms_eps, dtype, device)
            self.k_norm = RMSNorm(d, args.rms_eps, dtype, device)
        else:
            self.q_norm = self.k_norm = None
        self.rope = RoPE(d, args.rope_base, False)
        if args.sinks:
            full = torch.empty(args.n_heads, dtype=torch.float32, device=device).normal_(0.0, 1.0, generator=gen)
            self.register_buffer("sinks", full[rank * self.n_heads:(rank + 1) * self.n_heads].clone())
        else:
            self.sinks = None


class MLP(nn.Module):
    def __init__(self, args: ModelArgs, gen, dtype, device, rank: int, world: int) -> None:
        super().__init__()
        gate = _rand_qlinear(args.hidden, args.intermediate, args, gen, dtype, device).shard_rows(rank, world)
        up = _rand_qlinear(args.hidden, args.intermediate, args, gen, dtype, device).shard_rows(rank, world)
        self.gate_up_proj = _cat_rows([gate, up])
        self.down_proj = _rand_qlinear(args.intermediate, args.hidden, args, gen, dtype, device).shard_cols(rank, world)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self.gate_up_proj.swiglu:  # SiLU * up fused into the gate_up epilogue (streaming layout)
            f16 = getattr(self.down_proj, "weight_blob", None) is not None or \
                getattr(self.down_proj, "weight_tc", None) is not None
            return self.down_proj(self.gate_up_proj(x, out_f16=f16))
        return self.down_proj(get_ops().silu_mul(self.gate_up_proj(x)))


class MoEMLP(nn.Module):
    """Router + top-k quantized experts (gpt-oss / Qwen3-MoE wiring of mlx-lm); tensor parallel over the experts'
    intermediate dimension, so the output is this rank's partial sum like the dense MLP's."""

    def __init__(self, args: ModelArgs, gen, dtype, device, rank: int, world: int) -> None:
        super().__init__()
        from ..quant.switch_linear import MoEBlock, QuantizedSwitchLinear, Router, SwitchGLU

        e, h, i = args.num_experts, args.hidden, args.intermediate

        def rand(n, k, std):
            w = torch.empty((e, n, k), dtype=torch.float32, device=device).normal_(0.0, std, generator=gen).to(dtype)
            b = None
            if args.moe_bias:
                b = torch.empty((e, n), dtype=torch.float32, device=device).normal_(0.0, 0.02, generator=gen).to(dtype)
            return QuantizedSwitchLinear.from_dense(w, b, args.group_size, args.bits, dtype)

        glu = SwitchGLU(h, i // world, e, args.moe_bias, args.group_size, args.bits, dtype, device,
                        activation=args.moe_activation)
        glu.gate_proj = rand(i, h, 0.05).shard_rows(rank, world)
        glu.up_proj = rand(i, h, 0.05).shard_rows(rank, world)
        glu.down_proj = rand(h, i, 0.05).shard_cols(rank, world)
        router = Router(h, e, args.moe_bias, dtype, device)
        router.weight.copy_(torch.empty((e, h), dtype=torch.float32, device=device).normal_(0.0, 0.5, generator=gen))
        if router.bias is not None:
            router.bias.copy_(torch.empty((e,), dtype=torch.float32, device=device).normal_(0.0, 0.1, generator=gen))
        self.block = MoEBlock(router, glu, args.experts_per_tok, args.moe_routing)
        self.dtype = dtype

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.dtype != self.dtype:  # fp16 hand-over from the norm: the values are bf16-representable, the cast is exact
            x = x.to(self.dtype)
        return self.block(x)


class DecoderLayer(nn.Module):
    def __init__(self, args: ModelArgs, gen, dtype, device, rank: int, world: int) -> None:
        super().__init__()
        self.self_attn = Attention(args, gen, dtype, device, rank, world)
        self.input_layernorm = RMSNorm(args.hidden, args.rms_eps, dtype, device)
        if args.mlp and args.num_experts > 0:
            self.mlp = MoEMLP(args, gen, dtype, device, rank, world)
        else:
            self.mlp = MLP(args, gen, dtype, device, rank, world) if args.mlp else None
        self.post_attention_layernorm = RMSNorm(args.hidden, args.rms_eps, dtype, device) if args.mlp else None


class Model(nn.Module):
    """``forward(input_ids [T], ctx|meta, kv_cache) -> logits`` for the packed batch."""

    def __init__(self, args: ModelArgs, dtype=torch.bfloat16, device="cuda", seed: int = 0,
                 rank: int = 0, world: int = 1, allreduce=None, prepack: bool = True,
                 prefill_layout: bool = False, vocab_parallel: bool | None = None) -> None:
        super().__init__()
        self.args = args
        self.rank, self.world = rank, world
        self._allreduce = allreduce
        # Vocab-parallel head: each rank keeps vocab / world rows of the head and the greedy token comes
        # from a (max, id) exchange (``greedy``); ``forward`` then returns this rank's logits slice only.
        if vocab_parallel is None:
            vocab_parallel = world > 1 and hasattr(allreduce, "argmax_vocab_parallel") and args.vocab % world == 0
        if vocab_parallel and (world < 2 or args.vocab % world):
            raise ValueError(f"vocab-parallel head needs world > 1 and vocab ({args.vocab}) divisible by it ({world})")
        self.vocab_parallel = bool(vocab_parallel)
        self.vocab_offset = rank * (args.vocab // world) if self.vocab_parallel else 0
        gen = torch.Generator(device=device).manual_seed(seed)
        emb = torch.empty((args.vocab, args.hidden), dtype=torch.float32, device=device).normal_(0, 0.02, generator=gen)
        self.embed_tokens = QuantizedEmbedding.from_dense(emb.to(dtype), args.group_size, args.bits, dtype)
        del emb
        self.layers = nn.ModuleList(
            [DecoderLayer(args, gen, dtype, device, rank, world) for _ in range(args.num_layers)])
        self.norm = RMSNorm(args.hidden, args.rms_eps, dtype, device)
        if args.tie_embeddings:
            self.lm_head = None
        else:
            self.lm_head = _rand_qlinear(args.hidden, args.vocab, args, gen, dtype, device)
            if self.vocab_parallel:
                self.lm_head = self.lm_head.shard_rows(rank, world)
        self._f16_acts = False
        self._prefill_layout = prefill_layout
        if prepack:
            self.prepack()

    def prepack(self) -> None:
        """Load-time transform of every linear into the streaming layout; the MLX-layout
        ``weight`` is dropped (rebuilt on demand by ``mlx_weight()``) so the footprint is unchanged."""
        for m in self.modules():
            if isinstance(m, MLP):
                m.gate_up_proj.prepack(keep_mlx_layout=False, swiglu=True, prefill_layout=self._prefill_layout)
        for m in self.modules():
            if isinstance(m, QuantizedLinear) and getattr(m, "weight_blob", None) is None \
                    and getattr(m, "weight_tc", None) is None:
                m.prepack(keep_mlx_layout=False, prefill_layout=self._prefill_layout and m is not self.lm_head)
        if self.lm_head is None:
            n = self.args.vocab // self.world
            self.embed_tokens.prepack(rows=(self.rank * n, (self.rank + 1) * n) if self.vocab_parallel else None)
        # norms feed the stream-K linears fp16 directly (their operand type) when every consumer is one
        self._f16_acts = all(getattr(m, "weight_blob", None) is not None or getattr(m, "weight_tc", None) is not None
                             for m in self.modules() if isinstance(m, QuantizedLinear)) and (
            self.lm_head is not None or getattr(self.embed_tokens, "weight_blob", None) is not None)

    def _reduce(self, x: torch.Tensor) -> torch.Tensor:
        if self.world > 1:
            self._allreduce(x)
        return x

    def _sum_add_norm(self, norm: RMSNorm, part: torch.Tensor | None, h: torch.Tensor) -> torch.Tensor:
        """``h += sum over ranks(part); return norm(h)``: the tensor-parallel epilogue of a sub-block.  With the
        NVLink all-reduce object this is ONE kernel (all-reduce + residual add + RMSNorm, csrc/nvlink_allreduce.cu);
        otherwise all-reduce, then the fused add + norm."""
        if part is None:  # first layer: nothing to add yet
            return norm(h, out_f16=self._f16_acts)
        # push form first: the residual stream stays valid only on the rows this rank owns (r % world) -- nothing
        # but these sums reads it, and every sum of a forward pass takes the same branch (same shapes throughout)
        sharded = getattr(self._allreduce, "fused_add_rmsnorm_sharded", None) if self.world > 1 else None
        if sharded is not None:
            f16 = self._f16_acts and part.dtype == torch.bfloat16
            x = sharded(part, h, norm.weight, norm.eps, torch.float16 if f16 else None)
            if x is not None:
                return x
        fused = getattr(self._allreduce, "fused_add_rmsnorm", None) if self.world > 1 else None
        if fused is not None:
            f16 = self._f16_acts and part.dtype == torch.bfloat16
            x = fused(part, h, norm.weight, norm.eps, torch.float16 if f16 else None)
            if x is not None:
                return x
        return norm(self._reduce(part), residual=h, out_f16=self._f16_acts)

    def forward(self, input_ids: torch.Tensor, ctx, kv_cache, meta=None, logits_rows=None) -> torch.Tensor:
        h = self.embed_tokens(input_ids)  # residual stream [T, hidden]
        part = None                        # this rank's partial output of the previous sub-block (not yet summed)
        for i, layer in enumerate(self.layers):
            x = self._sum_add_norm(layer.input_layernorm, part, h)
            part = sdpa_forward(layer.self_attn, x, ctx, kv_cache, i, meta=meta)
            if layer.mlp is None:
                continue
            x = self._sum_add_norm(layer.post_attention_layernorm, part, h)  # h += a; x = norm(h)
            part = layer.mlp(x)
        x = self._sum_add_norm(self.norm, part, h)
        if logits_rows is not None:
            x = x.index_select(0, logits_rows)
        return self.embed_tokens.as_linear(x) if self.lm_head is None else self.lm_head(x)

    def greedy(self, logits: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
        """Greedy token per logits row (v1/sampling_batch.py:342-348: first maximum).  With a vocab-parallel head the
        rows are this rank's slice and the winner comes out of the NVLink (max, id) exchange, the same on every rank."""
        if self.vocab_parallel:
            return self._allreduce.argmax_vocab_parallel(logits, self.vocab_offset, out=out)
        return get_ops().argmax(logits, out=out)

    def weight_bytes(self) -> int:
        return sum(b.numel() * b.element_size() for b in self.buffers())
