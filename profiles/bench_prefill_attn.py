#!/usr/bin/env python
"""Timing of the tcgen05 tile kernel (csrc/paged_attention_prefill.cu) at the cfg-c attention shape.

    python profiles/bench_prefill_attn.py [--hq 32 --hkv 4 --d 128] [--chunk 2048] [--offsets 0,2048,6144]
                                          [--decodes 32] [--kernel 0|1|3]

One 2048-token prefill chunk continuing at `offset` + `decodes` decode rows at ctx U[512, 8192] (SURVEY 8(d) cfg c;
default head counts = Qwen3-32B at TP=2).  CUDA events over 10 launches after 3 warm-ups; flops =
4 * D * H_q * sum_i q_i * (ctx_i + (q_i + 1) / 2) (causal), against MEASURED_PEAKS.json bf16_tflops (burst).
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vllm_metal_b200.metal import get_ops  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--hq", type=int, default=32)
    ap.add_argument("--hkv", type=int, default=4)
    ap.add_argument("--d", type=int, default=128)
    ap.add_argument("--chunk", type=int, default=2048)
    ap.add_argument("--offsets", default="0,2048,6144")
    ap.add_argument("--decodes", type=int, default=32)
    ap.add_argument("--kernel", type=int, default=0)
    ap.add_argument("--window", type=int, default=-1)
    ap.add_argument("--bs", type=int, default=16, help="KV block size")
    args = ap.parse_args()
    peak = 1600.0
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peak = float(json.load(f)["bf16_tflops"])
    except Exception:  # noqa: BLE001
        pass
    ops = get_ops()
    bs = args.bs
    rng = np.random.default_rng(4)
    for off in [int(t) for t in args.offsets.split(",")]:
        dec = rng.integers(512, 8192, args.decodes).tolist()
        seqs = [(1, n + 1) for n in dec] + [(args.chunk, off + args.chunk)]
        per = [(kv + bs - 1) // bs for _, kv in seqs]
        nb = 1 + sum(per)
        g = torch.Generator(device="cuda").manual_seed(off)
        kc = torch.randn((nb, bs, args.hkv, args.d), device="cuda", generator=g).bfloat16()
        vc = torch.randn((nb, bs, args.hkv, args.d), device="cuda", generator=g).bfloat16()
        perm = (rng.permutation(nb - 1) + 1).astype(np.int32)
        bt = np.zeros((len(seqs), max(per)), np.int32)
        o = 0
        for i, n in enumerate(per):
            bt[i, :n] = perm[o:o + n]
            o += n
        t = sum(q for q, _ in seqs)
        q = torch.randn((t, args.hq, args.d), device="cuda", generator=g).bfloat16()
        out = torch.empty_like(q)
        cu = torch.tensor(np.concatenate([[0], np.cumsum([q_ for q_, _ in seqs])]).astype(np.int32), device="cuda")
        sl = torch.tensor([kv for _, kv in seqs], dtype=torch.int32, device="cuda")
        btd = torch.from_numpy(bt).cuda()
        msl = max(kv for _, kv in seqs)

        def run():
            ops.paged_attention_primitive(q, kc, vc, args.hkv, args.d ** -0.5, 0.0, btd, sl, cu, bs, msl, args.window, out,
                                          _kernel=args.kernel)

        for _ in range(3):
            run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            run()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        flops = 4.0 * args.d * args.hq * sum(ql * ((kv - ql) + (ql + 1) / 2) for ql, kv in seqs)
        pf_flops = 4.0 * args.d * args.hq * args.chunk * (off + (args.chunk + 1) / 2)
        print(f"chunk {args.chunk} @ offset {off:5d} + {args.decodes} decode rows, {args.hq}q/{args.hkv}kv x {args.d}, kernel={args.kernel} bs={bs}: "
              f"{ms * 1e3:8.1f} us  {flops / ms / 1e9:7.1f} TFLOP/s ({flops / ms / 1e9 / peak:.3f} of {peak:.0f} bf16 peak; "
              f"prefill chunk alone is {pf_flops / flops:.2f} of the flops)", flush=True)


if __name__ == "__main__":
    main()
