#!/usr/bin/env python
"""MoE expert block (quant/switch_linear.py) at the gpt-oss-20b shape: 32 experts, top-4, hidden = inter = 2880, int4 g64.

    python profiles/bench_moe.py [--tokens 32]
Eager timing with CUDA events (the block reads its per-expert row counts back once per call, so it is not graph
captured); bytes = the packed weights of the experts that were hit (each streamed once) + activations.
"""
import argparse, json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vllm_metal_b200.quant.switch_linear import MoEBlock, QuantizedSwitchLinear, Router, SwitchGLU  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--tokens", type=int, default=32)
a = ap.parse_args()
peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
H = I = 2880
E, K = 32, 4
g = torch.Generator(device="cuda").manual_seed(0)
glu = SwitchGLU(H, I, E, True, 64, 4, torch.bfloat16, "cuda", activation="gpt_oss")
for name, (n, k) in (("gate_proj", (I, H)), ("up_proj", (I, H)), ("down_proj", (H, I))):
    w = torch.randn((E, n, k), device="cuda", generator=g) * (0.7 / k ** 0.5)
    setattr(glu, name, QuantizedSwitchLinear.from_dense(w, torch.zeros((E, n), device="cuda"), 64, 4, torch.bfloat16))
router = Router(H, E, True, torch.bfloat16, "cuda")
router.weight.copy_(torch.randn((E, H), device="cuda", generator=g) * 0.5)
blk = MoEBlock(router, glu, K)
x = torch.randn((a.tokens, H), device="cuda", generator=g).bfloat16()
for _ in range(3):
    y = blk(x)
idx, _ = blk.route(x)
hit = int(torch.unique(idx).numel())
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    y = blk(x)
e1.record()
torch.cuda.synchronize()
us = e0.elapsed_time(e1) * 1e3 / 20
wbytes = hit * 3 * (H * I // 2 + 2 * H * I // 64 * 2)
print(f"tokens={a.tokens} experts hit {hit}/{E} top-{K}: {us:8.1f} us per block, {3 * hit} expert GEMM launches, "
      f"{wbytes / 1e6:.1f} MB of expert weights -> {wbytes / us / 1e3:.0f} GB/s ({wbytes / us / 1e3 / peak:.3f} of {peak:.0f})",
      flush=True)
