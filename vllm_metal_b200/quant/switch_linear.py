"""Quantized mixture-of-experts MLP: ``QuantizedSwitchLinear`` / ``SwitchGLU`` / ``MoEBlock``.

The reference reaches MoE models (gpt-oss: 32 experts, top-4) through mlx-lm's ``SwitchGLU`` whose expert
projections are ``QuantizedSwitchLinear`` modules -- named in vllm_metal/quant/awq_loader.py:171-173 -- i.e.
``mx.gather_qmm(x, weight, scales, biases, rhs_indices=indices, transpose=True, group_size, bits)`` over stacked
expert weights ``[E, N, K*bits/32]`` (third-party arithmetic: mlx 0.32.0 / mlx-lm 254d153, not under
/root/reference -> parity unpinned, semantics restated in oracle/moe.py).  Same attribute names and shapes here.

B200 form: the (token, expert) pairs are sorted by expert and every expert that received rows runs ONE affine
quantized GEMM over its gathered rows through the same kernels as the dense linears (`b200_qlinear_experts`: one
C-ABI call per projection that loops over the experts): each
expert's weights are streamed once per step however many tokens picked it.  The per-expert row counts are read back
to the host once per block (one small D2H copy): this is the functional path of row N2 -- launch-bound at decode
batch sizes (3 x experts-hit launches per layer); a grouped persistent kernel is the next step and is not built.
"""

from __future__ import annotations

import torch
from torch import nn

from ..metal import get_ops
from .affine import quantize


class QuantizedSwitchLinear(nn.Module):
    """Stacked expert linears: ``weight`` uint32 ``[E, N, K*bits/32]``, ``scales`` / ``biases`` ``[E, N, K/g]``,
    optional dense ``bias`` ``[E, N]`` (mlx-lm ``QuantizedSwitchLinear``)."""

    def __init__(self, in_features: int, out_features: int, num_experts: int, bias: bool = False,
                 group_size: int = 64, bits: int = 4, dtype: torch.dtype = torch.bfloat16, device=None) -> None:
        super().__init__()
        if in_features % group_size:
            raise ValueError(f"in_features {in_features} not divisible by group_size {group_size}")
        self.group_size, self.bits = group_size, bits
        e, n, k = num_experts, out_features, in_features
        self.register_buffer("weight", torch.zeros((e, n, k * bits // 32), dtype=torch.int32, device=device))
        self.register_buffer("scales", torch.zeros((e, n, k // group_size), dtype=dtype, device=device))
        self.register_buffer("biases", torch.zeros((e, n, k // group_size), dtype=dtype, device=device))
        if bias:
            self.register_buffer("bias", torch.zeros((e, n), dtype=dtype, device=device))
        else:
            self.bias = None

    @property
    def num_experts(self) -> int:
        return self.scales.shape[0]

    @property
    def out_features(self) -> int:
        return self.scales.shape[1]

    @property
    def in_features(self) -> int:
        return self.scales.shape[2] * self.group_size

    @classmethod
    def from_dense(cls, weight: torch.Tensor, bias: torch.Tensor | None = None, group_size: int = 64, bits: int = 4,
                   dtype: torch.dtype = torch.bfloat16) -> QuantizedSwitchLinear:
        """``weight`` ``[E, N, K]`` dense; every expert is quantized like a ``QuantizedLinear``."""
        e, n, k = weight.shape
        m = cls(k, n, e, bias is not None, group_size, bits, dtype, weight.device)
        for i in range(e):
            w, s, b = quantize(weight[i], group_size, bits, dtype)
            m.weight[i].copy_(w)
            m.scales[i].copy_(s)
            m.biases[i].copy_(b)
        if bias is not None:
            m.bias.copy_(bias.to(dtype))
        return m

    def _slice(self, rows: slice, groups: slice, words: slice, keep_bias: bool) -> QuantizedSwitchLinear:
        w, sc, bi = self.weight[:, rows, words], self.scales[:, rows, groups], self.biases[:, rows, groups]
        m = QuantizedSwitchLinear(sc.shape[2] * self.group_size, sc.shape[1], self.num_experts,
                                  self.bias is not None and keep_bias, self.group_size, self.bits, self.scales.dtype,
                                  self.weight.device)
        m.weight.copy_(w)
        m.scales.copy_(sc)
        m.biases.copy_(bi)
        if m.bias is not None:
            m.bias.copy_(self.bias[:, rows])
        return m

    def shard_rows(self, rank: int, world: int) -> QuantizedSwitchLinear:
        """Column-parallel: every expert's output rows [rank*N/world, ...) (gate / up of a tensor-parallel MoE)."""
        n = self.out_features
        if n % world:
            raise ValueError(f"out_features {n} not divisible by world size {world}")
        return self._slice(slice(rank * n // world, (rank + 1) * n // world), slice(None), slice(None), True)

    def shard_cols(self, rank: int, world: int) -> QuantizedSwitchLinear:
        """Row-parallel along K on group boundaries (down); the dense bias stays on rank 0."""
        k = self.in_features
        if k % world or (k // world) % self.group_size:
            raise ValueError(f"in_features {k} / {world} must be a multiple of group_size {self.group_size}")
        kw = k // world
        return self._slice(slice(None), slice(rank * kw // self.group_size, (rank + 1) * kw // self.group_size),
                           slice(rank * kw * self.bits // 32, (rank + 1) * kw * self.bits // 32), rank == 0)

    def forward_sorted(self, x_sorted: torch.Tensor, counts: list[int]) -> torch.Tensor:
        """``x_sorted`` ``[P, K]``: rows grouped by expert (``counts[e]`` consecutive rows for expert e).  One C-ABI call
        (``b200_qlinear_experts``) that launches one quantized GEMM per expert with rows; returns ``[P, N]``."""
        return get_ops().quantized_matmul_experts(x_sorted, self.weight, self.scales, self.biases, counts,
                                                  self.group_size, self.bits, bias=self.bias)

    def forward(self, x: torch.Tensor, indices: torch.Tensor) -> torch.Tensor:
        """``mx.gather_qmm`` form: ``x`` ``[T, K]`` (shared by a token's experts) or ``[T, k, K]``; ``indices``
        ``[T, k]`` -> ``[T, k, N]``."""
        t, k = indices.shape
        order, counts = _sort_pairs(indices, self.num_experts)
        rows = x if x.dim() == 3 else x[:, None, :].expand(t, k, x.shape[-1])
        xs = rows.reshape(t * k, -1).index_select(0, order)
        ys = self.forward_sorted(_f16_operand(xs.contiguous(), self), counts)
        out = torch.empty_like(ys)
        out.index_copy_(0, order, ys)
        return out.reshape(t, k, -1)


def _sort_pairs(indices: torch.Tensor, num_experts: int) -> tuple[torch.Tensor, list[int]]:
    flat = indices.reshape(-1).to(torch.int64)
    order = torch.argsort(flat, stable=True)
    counts = torch.bincount(flat, minlength=num_experts).tolist()  # the one host read-back of the block
    return order, counts


def _f16_operand(x: torch.Tensor, lin: QuantizedSwitchLinear) -> torch.Tensor:
    """The tensor-core linears multiply fp16 operands and would cast every expert's slice of a bf16 ``x`` on their own
    (one more launch per expert): cast the gathered rows once for all experts.  Exact: the values are bf16."""
    ops = get_ops()
    if (x.dtype == torch.bfloat16 and lin.scales.dtype == torch.bfloat16 and lin.in_features % 64 == 0
            and x.data_ptr() % 16 == 0 and x.stride(0) % 8 == 0 and getattr(ops, "_qlinear_kernel", 0) != 1):
        return ops.cast_f16(x)
    return x


def _act_silu(gate: torch.Tensor, up: torch.Tensor) -> torch.Tensor:
    return torch.nn.functional.silu(gate) * up


def _act_gpt_oss(gate: torch.Tensor, up: torch.Tensor, alpha: float = 1.702, limit: float = 7.0) -> torch.Tensor:
    """gpt-oss's clipped SwiGLU (mlx-lm gpt_oss.py ``swiglu``): ``g = min(gate, limit)``, ``u = clip(up, +-limit)``,
    ``g * sigmoid(alpha * g) * (u + 1)``."""
    g = gate.clamp(max=limit)
    u = up.clamp(min=-limit, max=limit)
    return g * torch.sigmoid(alpha * g) * (u + 1.0)


ACTIVATIONS = {"silu": _act_silu, "gpt_oss": _act_gpt_oss}


class SwitchGLU(nn.Module):
    """``down(act(gate(x), up(x)))`` per selected expert (mlx-lm ``SwitchGLU``).  The activation runs in fp32 on the
    dtype-rounded projections and is rounded once."""

    def __init__(self, hidden: int, inter: int, num_experts: int, bias: bool = False, group_size: int = 64,
                 bits: int = 4, dtype: torch.dtype = torch.bfloat16, device=None, activation: str = "silu") -> None:
        super().__init__()
        mk = lambda i, o: QuantizedSwitchLinear(i, o, num_experts, bias, group_size, bits, dtype, device)  # noqa: E731
        self.gate_proj, self.up_proj, self.down_proj = mk(hidden, inter), mk(hidden, inter), mk(inter, hidden)
        self.activation = activation

    def forward(self, x: torch.Tensor, indices: torch.Tensor) -> torch.Tensor:
        t, k = indices.shape
        order, counts = _sort_pairs(indices, self.gate_proj.num_experts)
        token_of_pair = torch.div(order, k, rounding_mode="floor")
        xs = x.index_select(0, token_of_pair).contiguous()  # rows grouped by expert: gathered once for gate and up
        xs = _f16_operand(xs, self.gate_proj)
        g = self.gate_proj.forward_sorted(xs, counts)
        u = self.up_proj.forward_sorted(xs, counts)
        h = ACTIVATIONS[self.activation](g.float(), u.float()).to(g.dtype)
        ys = self.down_proj.forward_sorted(_f16_operand(h, self.down_proj), counts)
        out = torch.empty_like(ys)
        out.index_copy_(0, order, ys)
        return out.reshape(t, k, -1)


class Router(nn.Module):
    """Dense router ``x @ W^T (+ b)`` (mlx-lm: ``nn.Linear(hidden, num_experts)``): fp32 accumulate, logits rounded to
    the activation dtype."""

    def __init__(self, hidden: int, num_experts: int, bias: bool, dtype: torch.dtype = torch.bfloat16, device=None) -> None:
        super().__init__()
        self.register_buffer("weight", torch.zeros((num_experts, hidden), dtype=dtype, device=device))
        if bias:
            self.register_buffer("bias", torch.zeros((num_experts,), dtype=dtype, device=device))
        else:
            self.bias = None

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        y = x.float() @ self.weight.float().t()
        if self.bias is not None:
            y = y + self.bias.float()
        return y.to(x.dtype)


class MoEBlock(nn.Module):
    """Router + top-k experts + weighted sum.

    ``routing="topk_softmax"`` (gpt-oss): top-k of the router logits, softmax over the k selected logits.
    ``routing="softmax_topk"`` (Qwen3-MoE style): softmax over all experts, top-k, optionally renormalised.
    Router and softmax in fp32; ``y = sum_j score_j * expert_j(x)`` accumulated in fp32, rounded once."""

    def __init__(self, router: nn.Module, experts: SwitchGLU, top_k: int, routing: str = "topk_softmax",
                 norm_topk_prob: bool = True) -> None:
        super().__init__()
        if routing not in ("topk_softmax", "softmax_topk"):
            raise ValueError(f"unknown routing {routing!r}")
        self.router, self.experts, self.top_k = router, experts, int(top_k)
        self.routing, self.norm_topk_prob = routing, norm_topk_prob
        self.trace: list | None = None  # tests: a list collects (indices, scores) of every forward

    def route(self, x: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor]:
        logits = self.router(x).float()
        # stable descending sort: equal scores (bf16-rounded logits do tie) go to the lowest expert id, on every rank
        # and in the oracle alike -- torch.topk leaves the order of ties unspecified
        if self.routing == "topk_softmax":
            vals, idx = torch.sort(logits, dim=-1, descending=True, stable=True)
            return idx[:, :self.top_k], torch.softmax(vals[:, :self.top_k], dim=-1)
        probs = torch.softmax(logits, dim=-1)
        vals, idx = torch.sort(probs, dim=-1, descending=True, stable=True)
        vals, idx = vals[:, :self.top_k], idx[:, :self.top_k]
        if self.norm_topk_prob:
            vals = vals / vals.sum(dim=-1, keepdim=True)
        return idx, vals

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        lead = x.shape[:-1]
        x2 = x.reshape(-1, x.shape[-1])
        idx, scores = self.route(x2)
        if self.trace is not None:
            self.trace.append((idx.cpu().numpy(), scores.cpu().numpy()))
        y = self.experts(x2, idx)  # [T, k, hidden]
        out = (y.float() * scores[..., None]).sum(dim=1).to(x.dtype)
        return out.reshape(*lead, -1)
