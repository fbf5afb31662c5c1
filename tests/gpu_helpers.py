"""Helpers shared by the ``-m gpu`` parity tests (never imported by the product)."""

from __future__ import annotations

import numpy as np
import torch

from oracle import paged_attention as oattn

TORCH_DT = {"float32": torch.float32, "float16": torch.float16, "bfloat16": torch.bfloat16}

# Tolerances: BASELINE.json north_star for bf16 (2e-2 max-abs / 1e-2 relative); the
# reference's own tables for the others (tests/test_split_kv_decode.py:46-50).
TOL = {"bfloat16": (2e-2, 1e-2), "float16": (1.5e-2, 1e-2), "float32": (1e-3, 1e-3)}


def to_np(t: torch.Tensor) -> np.ndarray:
    return t.detach().float().cpu().numpy()


def make_paged_case(seqs, hq, hkv, d, block_size, dtype="bfloat16", seed=0, device="cuda",
                    extra_blocks=3, sinks=False):
    """``seqs``: list of (q_len, kv_len).  Random cache pool with a random
    (non-contiguous) page permutation; block 0 is reserved as the padding block
    like tools/nax_prefill_parity.py:114."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    tdt = TORCH_DT[dtype]
    blocks_per_seq = [(kv + block_size - 1) // block_size for _, kv in seqs]
    nb = 1 + sum(blocks_per_seq) + extra_blocks
    kc = torch.randn((nb, block_size, hkv, d), generator=g).to(tdt)
    vc = torch.randn((nb, block_size, hkv, d), generator=g).to(tdt)
    perm = (torch.randperm(nb - 1, generator=g) + 1).tolist()
    maxb = max(blocks_per_seq)
    bt = np.zeros((len(seqs), maxb), np.int32)
    off = 0
    for i, n in enumerate(blocks_per_seq):
        bt[i, :n] = perm[off:off + n]
        off += n
    t_total = sum(q for q, _ in seqs)
    q = torch.randn((t_total, hq, d), generator=g).to(tdt)
    cu = np.concatenate([[0], np.cumsum([q_ for q_, _ in seqs])]).astype(np.int32)
    seq_lens = np.asarray([kv for _, kv in seqs], np.int32)
    sk = torch.randn(hq, generator=g) * 2.0 if sinks else None
    dev = lambda x: x.to(device)  # noqa: E731
    return dict(
        q=dev(q), kc=dev(kc), vc=dev(vc), bt=torch.from_numpy(bt).to(device),
        seq_lens=torch.from_numpy(seq_lens).to(device), cu=torch.from_numpy(cu).to(device),
        sinks=None if sk is None else dev(sk.float()), hq=hq, hkv=hkv, d=d, block_size=block_size,
        dtype=dtype, scale=float(d) ** -0.5, max_seq_len=int(seq_lens.max()),
        np=dict(q=to_np(q), kc=to_np(kc), vc=to_np(vc), bt=bt, seq_lens=seq_lens, cu=cu,
                sinks=None if sk is None else sk.float().numpy()),
    )


def oracle_attention(c, window=-1, softcap=0.0, out_init=None):
    n = c["np"]
    return oattn.paged_attention_kernel_exact(
        n["q"], n["kc"], n["vc"], n["bt"], n["seq_lens"], n["cu"], c["scale"], window, softcap,
        n["sinks"], out=out_init, dtype=c["dtype"])


def run_attention(ops, c, window=-1, softcap=0.0, force_partitions=0, kernel=0, out=None):
    if out is None:
        out = torch.zeros_like(c["q"])
    ops.paged_attention_primitive(
        c["q"], c["kc"], c["vc"], c["hkv"], c["scale"], softcap, c["bt"], c["seq_lens"], c["cu"],
        c["block_size"], c["max_seq_len"], window, out, sinks=c["sinks"],
        _force_partitions=force_partitions, _kernel=kernel)
    torch.cuda.synchronize()
    return out


def assert_close(got: torch.Tensor | np.ndarray, want: np.ndarray, dtype: str, what: str = ""):
    g = to_np(got) if isinstance(got, torch.Tensor) else got
    atol, rtol = TOL[dtype]
    assert np.isfinite(g).all(), f"{what}: non-finite output"
    err = np.abs(g - want)
    bound = atol + rtol * np.abs(want)
    if not (err <= bound).all():
        i = np.unravel_index((err - bound).argmax(), err.shape)
        raise AssertionError(
            f"{what}: max|err|={err.max():.4g} at {i} got={g[i]:.6g} want={want[i]:.6g} "
            f"(atol={atol}, rtol={rtol})")


# ---------------------------------------------------------------- model level ---

def moe_weights(block):
    """Host copies of a product ``MoEBlock`` for ``oracle.moe.moe_block``.  The product's routing of every forward is
    handed to the oracle (``forced``): expert choice is discrete, so a 1-ulp difference in the hidden state that
    reaches the router can flip an expert and change the output far beyond any tolerance; routing parity itself is
    tested on identical inputs in tests/test_gpu_moe.py and tests/test_host_logic.py."""
    block.trace = []
    def stack(m):
        return (m.weight.cpu().numpy().view(np.uint32), to_np(m.scales), to_np(m.biases),
                None if m.bias is None else to_np(m.bias))

    g = block.experts
    return dict(router_w=to_np(block.router.weight), router_b=None if block.router.bias is None else to_np(block.router.bias),
                gate=stack(g.gate_proj), up=stack(g.up_proj), down=stack(g.down_proj), top_k=block.top_k,
                group_size=g.gate_proj.group_size, bits=g.gate_proj.bits, routing=block.routing, act=g.activation,
                norm_topk_prob=block.norm_topk_prob, forced=block.trace)


def oracle_from_model(model, num_blocks, block_size, c_attention=False):
    """Build an ``oracle.decoder.OracleDecoder`` holding the SAME quantised weights
    as a product ``lm.decoder.Model`` (copied to host; nothing is recomputed)."""
    from oracle.decoder import OracleDecoder, QWeight

    def qw(m):
        w = m.mlx_weight() if hasattr(m, "mlx_weight") else m.weight  # MLX layout (undo the load-time prepack)
        return QWeight(w.cpu().numpy().view(np.uint32), to_np(m.scales), to_np(m.biases),
                       m.group_size, m.bits)

    a = model.args
    windows = a.sliding_window_per_layer()
    layers = []
    for i, layer in enumerate(model.layers):
        at = layer.self_attn
        layers.append(dict(
            input_norm=to_np(layer.input_layernorm.weight),
            post_norm=None if layer.post_attention_layernorm is None else to_np(layer.post_attention_layernorm.weight),
            qkv=qw(at.qkv_proj), o=qw(at.o_proj),
            q_norm=None if at.q_norm is None else to_np(at.q_norm.weight),
            k_norm=None if at.k_norm is None else to_np(at.k_norm.weight),
            gate_up=None if layer.mlp is None or hasattr(layer.mlp, "block") else qw(layer.mlp.gate_up_proj),
            down=None if layer.mlp is None or hasattr(layer.mlp, "block") else qw(layer.mlp.down_proj),
            moe=None if not hasattr(layer.mlp, "block") else moe_weights(layer.mlp.block),
            sinks=None if at.sinks is None else to_np(at.sinks), window=windows[i]))
    weights = dict(embed=qw(model.embed_tokens), lm_head=None if model.lm_head is None else qw(model.lm_head),
                   norm=to_np(model.norm.weight), layers=layers)
    return OracleDecoder(weights, a.n_heads, a.n_kv_heads, a.head_dim, a.rms_eps, a.rope_base, num_blocks,
                         block_size, "bfloat16", c_attention=c_attention)
