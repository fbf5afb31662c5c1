#!/usr/bin/env python
"""Spec-verify paged attention: K+1-row windows dispatched expanded (one row per segment) vs in window mode.

    python profiles/bench_spec_window.py [--bs 64 --ctx 4096 --hq 32 --hkv 8 --d 128 --layers 16 --window 4]
CUDA graph of `layers` launches over distinct caches (>> L2), CUDA events.  Bytes = the algorithmic KV bytes of the
step (every sequence's KV once per kv head) + q/out; the expanded dispatch streams them `window` times.
"""
import argparse, json, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vllm_metal_b200.metal import get_ops  # noqa: E402

ap = argparse.ArgumentParser()
for k, v in (("bs", 64), ("ctx", 4096), ("hq", 32), ("hkv", 8), ("d", 128), ("layers", 16), ("window", 4)):
    ap.add_argument("--" + k, type=int, default=v)
a = ap.parse_args()
ops = get_ops()
peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
blk, w = 16, a.window
per = (a.ctx + w + blk - 1) // blk + 1
nb = 1 + a.bs * per
g = torch.Generator(device="cuda").manual_seed(0)
kc = [torch.randn((nb, blk, a.hkv, a.d), device="cuda", generator=g).bfloat16() for _ in range(a.layers)]
vc = [torch.randn((nb, blk, a.hkv, a.d), device="cuda", generator=g).bfloat16() for _ in range(a.layers)]
bt = torch.from_numpy((np.random.default_rng(2).permutation(nb - 1) + 1).astype(np.int32).reshape(a.bs, per)).cuda()
q = torch.randn((a.bs * w, a.hq, a.d), device="cuda", generator=g).bfloat16()
layouts = {
    "expanded": (bt.repeat_interleave(w, 0).contiguous(),
                 (a.ctx + 1 + torch.arange(w, dtype=torch.int32, device="cuda")).repeat(a.bs),
                 torch.arange(a.bs * w + 1, dtype=torch.int32, device="cuda"), 1),
    "windowed": (bt, torch.full((a.bs,), a.ctx + w, dtype=torch.int32, device="cuda"),
                 torch.arange(0, a.bs * w + 1, w, dtype=torch.int32, device="cuda"), w),
}
nbytes = a.bs * (2 * (a.ctx + w) * a.hkv * a.d * 2 + 2 * w * a.hq * a.d * 2 + per * 4)
outs = {}
for name, (tab, sl, cu, hint) in layouts.items():
    out = torch.empty_like(q)

    def run():
        for i in range(a.layers):
            ops.paged_attention_primitive(q, kc[i], vc[i], a.hkv, a.d ** -0.5, 0.0, tab, sl, cu, blk, a.ctx + w, -1, out,
                                          window_seqlen_q=hint)
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
    outs[name] = out.clone()
    print(f"bs={a.bs} ctx={a.ctx} {a.hq}q/{a.hkv}kv x{a.d} window={w} {name:9s}: {us:7.1f} us/layer  "
          f"{nbytes / us / 1e3:6.0f} GB/s algorithmic ({nbytes / us / 1e3 / peak:.3f} of {peak:.0f})", flush=True)
print("bitwise equal:", bool(torch.equal(outs["expanded"], outs["windowed"])), flush=True)
