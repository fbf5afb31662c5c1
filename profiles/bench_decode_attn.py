#!/usr/bin/env python
"""Paged decode attention alone at a BASELINE decode shape: heuristic split-KV plan vs single pass.

    python profiles/bench_decode_attn.py [--bs 64 --ctx 4096 --hq 32 --hkv 8 --d 128 --layers 32]
CUDA graph of `layers` launches over distinct caches (>> L2), CUDA events; GB/s on the algorithmic bytes.
"""
import argparse, json, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vllm_metal_b200.metal import get_ops  # noqa: E402

ap = argparse.ArgumentParser()
for k, v in (("bs", 64), ("ctx", 4096), ("hq", 32), ("hkv", 8), ("d", 128), ("layers", 32)):
    ap.add_argument("--" + k, type=int, default=v)
a = ap.parse_args()
ops = get_ops()
peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
blk = 16
per = (a.ctx + blk) // blk + 1
nb = 1 + a.bs * per
g = torch.Generator(device="cuda").manual_seed(0)
kc = [torch.randn((nb, blk, a.hkv, a.d), device="cuda", generator=g).bfloat16() for _ in range(a.layers)]
vc = [torch.randn((nb, blk, a.hkv, a.d), device="cuda", generator=g).bfloat16() for _ in range(a.layers)]
bt = torch.from_numpy((np.random.default_rng(2).permutation(nb - 1) + 1).astype(np.int32).reshape(a.bs, per)).cuda()
sl = torch.full((a.bs,), a.ctx + 1, dtype=torch.int32, device="cuda")
cu = torch.arange(a.bs + 1, dtype=torch.int32, device="cuda")
q = torch.randn((a.bs, a.hq, a.d), device="cuda", generator=g).bfloat16()
out = torch.empty_like(q)
nbytes = a.bs * (2 * (a.ctx + 1) * a.hkv * a.d * 2 + 2 * a.hq * a.d * 2 + per * 4)
for fp in (0, 1):
    def run():
        for i in range(a.layers):
            ops.paged_attention_primitive(q, kc[i], vc[i], a.hkv, a.d ** -0.5, 0.0, bt, sl, cu, blk, a.ctx + 1, -1, out,
                                          _force_partitions=fp)
    run(); torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        run()
    gr.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        gr.replay()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (5 * a.layers)
    print(f"bs={a.bs} ctx={a.ctx} {a.hq}q/{a.hkv}kv x{a.d} force_partitions={fp} ({'heuristic plan' if fp == 0 else 'single pass'}): "
          f"{us:7.1f} us/layer  {nbytes / us / 1e3:6.0f} GB/s ({nbytes / us / 1e3 / peak:.3f} of {peak:.0f})", flush=True)
