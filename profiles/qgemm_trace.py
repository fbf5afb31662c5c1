#!/usr/bin/env python
"""Per-CTA phase timeline of the tcgen05 quantized GEMM (b200_set_qlinear_trace).

    python profiles/qgemm_trace.py [gate_up,o_proj,...] [ksplit]

Each CTA stamps %globaltimer at: 0 start, 1 setup done, 2 previous kernel complete (PDL wait),
3 first packed weights landed, 4 dequant done, 5 accumulator complete, 6 epilogue done.
Prints, relative to the earliest CTA start of the launch, the median / max of every stamp and the
distribution of CTA start times (wave structure).  Tuning evidence only.
"""

from __future__ import annotations

import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from profiles.bench_kernels import SHAPES, make_qweights  # noqa: E402
from vllm_metal_b200.metal import get_ops  # noqa: E402

NAMES = ["start", "setup", "dep_met", "first_w", "dequant_done", "acc_done", "epi_done"]


def main():
    shapes = (sys.argv[1] if len(sys.argv) > 1 else "gate_up,o_proj,down,qkv").split(",")
    ks = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    ops = get_ops()
    gen = torch.Generator(device="cuda").manual_seed(0)
    m = int(os.environ.get("M", "64"))
    for name in shapes:
        n, k = SHAPES[name]
        ws = make_qweights(n, k, 4, 64, 4, gen)
        wps = [(ops.qlinear_prepack(w, 4), s, b) for (w, s, b) in ws]
        x = torch.randn((m, k), device="cuda", generator=gen).bfloat16()
        y = torch.empty((m, n), device="cuda", dtype=torch.bfloat16)
        ops.set_qlinear_tuning(ks, 0)
        trace = torch.zeros((8192, 64), dtype=torch.int64, device="cuda")
        for i in range(3):  # warm
            ops.quantized_matmul(x, *wps[i], group_size=64, bits=4, out=y, prepacked=True)
        torch.cuda.synchronize()
        # a dependent chain like the decode step: a small kernel writes x, then the GEMM
        x.add_(0.0)
        ops.set_qlinear_trace(trace)
        ops.quantized_matmul(x, *wps[3], group_size=64, bits=4, out=y, prepacked=True)
        ops.set_qlinear_trace(None)
        torch.cuda.synchronize()
        t = trace.cpu()
        t = t[t[:, 0] > 0]
        t0 = int(t[:, 0].min())
        rel = (t[:, :7] - t0).double() / 1e3  # us
        print(f"== {name} [{n} x {k}] M={m} ksplit={ks or 'auto'}: {t.shape[0]} CTAs, "
              f"span {float(rel[:, 6].max()):.1f} us")
        for j, nm in enumerate(NAMES):
            col = rel[:, j]
            col = col[col >= 0]
            if col.numel() == 0:
                continue
            print(f"   {nm:13s} min {float(col.min()):7.2f}  med {float(col.median()):7.2f}  max {float(col.max()):7.2f}")
        life = rel[:, 6] - rel[:, 0]
        fill = rel[:, 3] - rel[:, 0]
        deq = rel[:, 4] - rel[:, 3]
        tail = rel[:, 6] - rel[:, 4]
        print(f"   per CTA: life med {float(life.median()):.2f} us | start->first_w med {float(fill.median()):.2f} | "
              f"first_w->dequant_done med {float(deq.median()):.2f} max {float(deq.max()):.2f} | "
              f"dequant_done->end med {float(tail.median()):.2f}")
        cyc = t[:, 8:12].double()
        print(f"   dequant thread 0 cycles (median over CTAs): loop {float(cyc[:, 0].median()):.0f} | wait full_p "
              f"{float(cyc[:, 1].median()):.0f} | wait empty_ab {float(cyc[:, 2].median()):.0f} | scale/bias "
              f"{float(cyc[:, 3].median()):.0f}")
        c0 = int(torch.argmin(t[:, 0]))
        seen = [(int(t[c0, 16 + i]) - t0) / 1e3 for i in range(16) if int(t[c0, 16 + i]) > 0]
        issued = [(int(t[c0, 32 + i]) - t0) / 1e3 for i in range(16) if int(t[c0, 32 + i]) > 0]
        print("   CTA0 packed stage issued at: " + " ".join(f"{v:.2f}" for v in issued))
        print("   CTA0 packed stage seen   at: " + " ".join(f"{v:.2f}" for v in seen))
        starts = rel[:, 0].sort().values
        q = [float(starts[int(p * (len(starts) - 1))]) for p in (0.0, 0.25, 0.5, 0.66, 0.75, 0.9, 1.0)]
        print("   CTA start quantiles (0,25,50,66,75,90,100%): " + " ".join(f"{v:.1f}" for v in q))


if __name__ == "__main__":
    main()
