#!/usr/bin/env python
"""Tiny driver for `ncu --set full`: a few launches of one hot kernel at the headline shape.

    python profiles/ncu_target.py qgemm [gate_up|qkv|o_proj|down] [ksplit] [dequant_warps]
    python profiles/ncu_target.py attn
    WINDOW=4 python profiles/ncu_target.py window     # decode attention in spec-verify window mode

Weights / KV rotate over buffers larger than L2 so the profiled launch streams from HBM.
"""

from __future__ import annotations

import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from vllm_metal_b200.metal import get_ops  # noqa: E402


def main():
    what = sys.argv[1] if len(sys.argv) > 1 else "qgemm"
    ops = get_ops()
    gen = torch.Generator(device="cuda").manual_seed(0)
    if what == "qgemm":
        from profiles.bench_kernels import SHAPES, make_qweights

        name = sys.argv[2] if len(sys.argv) > 2 else "gate_up"
        ks = int(sys.argv[3]) if len(sys.argv) > 3 else 0
        dq = int(sys.argv[4]) if len(sys.argv) > 4 else 0
        m = int(os.environ.get("M", "64"))
        n, k = SHAPES[name]
        copies = max(3, (200 << 20) // (n * k // 2) + 1)
        ws = make_qweights(n, k, 4, 64, copies, gen)
        x = torch.randn((m, k), device="cuda", generator=gen).bfloat16()
        y = torch.empty((m, n), device="cuda", dtype=torch.bfloat16)
        ops.set_qlinear_tuning(ks, dq)
        pre = os.environ.get("PRE", "1") == "1"
        if pre:
            ws = [(ops.qlinear_prepack(w, 4), s, b) for (w, s, b) in ws]
        for _ in range(2):
            for w, s, b in ws:
                ops.quantized_matmul(x, w, s, b, group_size=64, bits=4, out=y, prepacked=pre)
        torch.cuda.synchronize()
        print("qgemm ok", name, float(y.float().abs().mean()))
    else:
        hq, hkv, d, bs, ctx, batch = 32, 8, 128, 16, 4096, 64
        blocks = ctx // bs + 1
        nb = 1 + batch * blocks
        layers = 3
        caches = [(torch.randn((nb, bs, hkv, d), device="cuda", generator=gen).bfloat16(),
                   torch.randn((nb, bs, hkv, d), device="cuda", generator=gen).bfloat16()) for _ in range(layers)]
        bt = (torch.randperm(nb - 1, device="cuda", generator=gen) + 1).int().reshape(batch, blocks)
        seq = torch.full((batch,), ctx + 1, dtype=torch.int32, device="cuda")
        cu = torch.arange(batch + 1, dtype=torch.int32, device="cuda")
        win = int(os.environ.get("WINDOW", "4")) if what == "window" else 1  # spec-verify window mode: K+1 rows per sequence
        if win > 1:
            seq = torch.full((batch,), ctx + 1, dtype=torch.int32, device="cuda")  # the window's rows end at ctx + 1
            cu = torch.arange(0, batch * win + 1, win, dtype=torch.int32, device="cuda")
        q = torch.randn((batch * win, hq, d), device="cuda", generator=gen).bfloat16()
        out = torch.empty_like(q)
        for _ in range(2):
            for kc, vc in caches:
                ops.paged_attention_primitive(q, kc, vc, hkv, d ** -0.5, 0.0, bt, seq, cu, bs, ctx + 1, -1, out,
                                              window_seqlen_q=win)
        torch.cuda.synchronize()
        print("attn ok", float(out.float().abs().mean()))


if __name__ == "__main__":
    main()
