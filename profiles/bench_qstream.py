#!/usr/bin/env python
"""Timing of the persistent stream-K quantized GEMM (csrc/qgemm_stream.cu) at the decode shapes.

    python profiles/bench_qstream.py [--m 64] [--model llama-8b|llama-8b-tp8|qwen3-32b|llama-70b-tp8]
                                     [--grids 0,148,74] [--trace] [--out file.json]

CUDA events over a rotation of distinct weight blobs larger than L2 (126 MB), captured in a CUDA graph
so host launch overhead does not pace the kernels.  Prints achieved GB/s on the algorithmic bytes
N*K*(bits/8 + 4/g) + (M*K + M*N_out)*2 against MEASURED_PEAKS.json and, with --legacy, the round-1
kernel (qgemm_tc.cu, MLX layout) beside it.  Tuning evidence; the judged numbers are bench.py's.
"""

from __future__ import annotations

import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from vllm_metal_b200.metal import get_ops  # noqa: E402

PEAK = 6458.0
try:
    with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
        PEAK = float(json.load(f)["hbm_gbs"])
except Exception:  # noqa: BLE001
    pass

MODELS = {  # name: (bits, {shape: (N, K, swiglu)})
    "llama-8b": (4, {"qkv": (6144, 4096, 0), "o_proj": (4096, 4096, 0), "gate_up": (28672, 4096, 1),
                     "down": (4096, 14336, 0), "lm_head": (128256, 4096, 0)}),
    "llama-8b-tp8": (4, {"qkv": (768, 4096, 0), "o_proj": (4096, 512, 0), "gate_up": (3584, 4096, 1),
                         "down": (4096, 1792, 0), "lm_head_vp": (16032, 4096, 0)}),
    "llama-8b-tp2": (4, {"qkv": (3072, 4096, 0), "o_proj": (4096, 2048, 0), "gate_up": (14336, 4096, 1),
                         "down": (4096, 7168, 0)}),
    "llama-8b-tp4": (4, {"qkv": (1536, 4096, 0), "o_proj": (4096, 1024, 0), "gate_up": (7168, 4096, 1),
                         "down": (4096, 3584, 0)}),
    "qwen3-32b": (8, {"qkv": (10240, 5120, 0), "o_proj": (5120, 8192, 0), "gate_up": (51200, 5120, 1),
                      "down": (5120, 25600, 0)}),
    "llama-70b-tp8": (4, {"qkv": (1280, 8192, 0), "o_proj": (8192, 1024, 0), "gate_up": (7168, 8192, 1),
                          "down": (8192, 3584, 0)}),
    "qwen3-0.6b": (4, {"qkv": (4096, 1024, 0), "o_proj": (1024, 2048, 0), "gate_up": (6144, 1024, 1),
                       "down": (1024, 3072, 0), "lm_head": (151936, 1024, 0)}),
}


def time_us(fn, n_rot, reps=5, warm=2):
    for _ in range(warm):
        for i in range(n_rot):
            fn(i)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(n_rot):
            fn(i)
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (reps * n_rot)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--m", type=int, default=64)
    ap.add_argument("--model", default="llama-8b")
    ap.add_argument("--group", type=int, default=64)
    ap.add_argument("--grids", default="0")
    ap.add_argument("--shapes", default="")
    ap.add_argument("--legacy", action="store_true")
    ap.add_argument("--trace", action="store_true")
    ap.add_argument("--no-pdl", action="store_true")
    ap.add_argument("--x16", action="store_true", help="fp16 x (no staging cast launch)")
    ap.add_argument("--units", type=int, default=0, help="with --trace: print CTA 0's first N per-unit pipeline stamps")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    ops = get_ops()
    if args.no_pdl:
        ops.set_pdl(False)
    gen = torch.Generator(device="cuda").manual_seed(0)
    bits, shapes = MODELS[args.model]
    m, g = args.m, args.group
    res = {"peak_gbs": PEAK, "m": m, "model": args.model, "shapes": {}}
    print(f"# {args.model} int{bits} g{g} M={m}; HBM peak {PEAK:.0f} GB/s (measured copy)")
    for name, (n, k, swiglu) in shapes.items():
        if args.shapes and name not in args.shapes.split(","):
            continue
        per = n * k * bits // 8
        copies = max(2, min(24, (300 << 20) // per + 1))
        blobs, mlx = [], []
        for _ in range(copies):
            w = torch.randint(-2**31, 2**31 - 1, (n, k * bits // 32), dtype=torch.int64, device="cuda", generator=gen).to(torch.int32)
            s = (torch.rand((n, k // g), device="cuda", generator=gen) * 0.01 + 0.001).bfloat16()
            b = (-8.0 * s.float() + 0.001 * torch.randn((n, k // g), device="cuda", generator=gen)).bfloat16()
            blobs.append(ops.qstream_prepack(w, s, b, g, bits, bool(swiglu)))
            if args.legacy and len(mlx) < copies:
                mlx.append((w, s, b))
        x = torch.randn((m, k), device="cuda", generator=gen).bfloat16()
        if args.x16:
            x = x.half()
        n_out = n // 2 if swiglu else n
        y = torch.empty((m, n_out), device="cuda", dtype=torch.bfloat16)
        nbytes = n * k * (bits / 8 + 4 / g) + m * (k + n_out) * 2
        row = {"N": n, "K": k, "swiglu": swiglu, "bytes": nbytes, "roofline_us": nbytes / PEAK / 1e3}
        line = f"{name:10s} [{n:6d} x {k:5d}]{' swiglu' if swiglu else '':7s} {nbytes / 1e6:7.1f} MB  roofline {nbytes / PEAK / 1e3:6.2f} us |"
        for grid in [int(t) for t in args.grids.split(",")]:
            ops.set_qstream_grid(grid)
            us = time_us(lambda i: ops.qstream_linear(x, blobs[i], n, k, g, bits, swiglu=bool(swiglu), out=y), copies)
            row[f"grid{grid}_us"] = us
            line += f" grid {grid:3d}: {us:7.2f} us {nbytes / us / 1e3:6.0f} GB/s ({nbytes / us / 1e3 / PEAK:4.2f}) |"
        ops.set_qstream_grid(0)
        if args.legacy:
            yl = torch.empty((m, n), device="cuda", dtype=torch.bfloat16)
            us = time_us(lambda i: ops.quantized_matmul(x, *mlx[i], group_size=g, bits=bits, out=yl), copies)  # x fp16 with --x16: no cast launch
            row["legacy_us"] = us
            line += f" round-1 kernel: {us:7.2f} us ({(n * k * (bits / 8 + 4 / g) + m * (k + n) * 2) / us / 1e3 / PEAK:4.2f})"
        if args.trace:
            tr = torch.zeros((148 * 16 + 64 * 8,), dtype=torch.int64, device="cuda")
            ops.set_qstream_trace(tr)
            flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            ops.qstream_linear(x, blobs[0], n, k, g, bits, swiglu=bool(swiglu), out=y)
            e1.record()
            torch.cuda.synchronize()
            line += f"\n             single cold launch (events, incl. cast): {e0.elapsed_time(e1) * 1e3:.2f} us"
            del flush
            ops.set_qstream_trace(None)
            units = tr[148 * 16:].view(64, 8).cpu()
            t = tr[:148 * 16].view(148, 16).cpu()
            act = t[:, 0] > 0
            t0 = t[act, 0].min()
            rel = (t[act][:, :14] - t0).float() / 1e3
            line += (f"\n             trace (us, mean/max over {int(act.sum())} CTAs): start {rel[:, 0].mean():.2f}/{rel[:, 0].max():.2f}"
                     f" setup {rel[:, 1].mean():.2f}/{rel[:, 1].max():.2f} x-ready {rel[:, 2].mean():.2f}/{rel[:, 2].max():.2f}"
                     f" first-w {rel[:, 5].mean():.2f}/{rel[:, 5].max():.2f} epi-pdl {rel[:, 9].mean():.2f}/{rel[:, 9].max():.2f}"
                     f" dequant-done {rel[:, 3].mean():.2f}/{rel[:, 3].max():.2f} mma-done {rel[:, 6].mean():.2f}/{rel[:, 6].max():.2f}"
                     f" first-acc {rel[:, 7].mean():.2f}/{rel[:, 7].max():.2f} epi-done {rel[:, 8].mean():.2f}/{rel[:, 8].max():.2f}"
                     f" peers-in {rel[:, 11].mean():.2f}/{rel[:, 11].max():.2f} fetched0 {rel[:, 12].mean():.2f}/{rel[:, 12].max():.2f}"
                     f" fetched1 {rel[:, 13].mean():.2f}/{rel[:, 13].max():.2f} finish-done {rel[:, 10].mean():.2f}/{rel[:, 10].max():.2f}")
        print(line, flush=True)
        if args.trace and args.units:
            c0 = int(units[0, 0])
            print("             CTA 0 per-unit SM clocks since its first w_full: w_full | math done | a_empty | STTM+arrive || a_full(mma) | committed")
            for i in range(min(args.units, 64)):
                u = units[i]
                if int(u[0]) == 0:
                    break
                print(f"             unit {i:2d}: {int(u[0]) - c0:7d} {int(u[1]) - c0:7d} {int(u[2]) - c0:7d} {int(u[5]) - c0:7d} || {int(u[3]) - c0:7d} {int(u[4]) - c0:7d}")
        res["shapes"][name] = row
        del blobs, mlx
        torch.cuda.empty_cache()
    if args.out:
        with open(args.out, "w") as f:
            json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
