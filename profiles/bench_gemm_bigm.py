#!/usr/bin/env python
"""Tensor-core quantized GEMM (csrc/qgemm_tc.cu) at M >= 256 on the BASELINE layer shapes.

    python profiles/bench_gemm_bigm.py [--ms 256,1024,2048,4096]
x in fp16 (what the norm / SwiGLU producers hand over), weights in the kernel's prepacked layout, CUDA graph of
several launches over distinct weight copies, CUDA events; NVML SM clock sampled during the timed region.
TFLOP/s = 2*M*N*K / t against MEASURED_PEAKS.json bf16_tflops (burst: the kernel is timed alone).
"""
import argparse, json, os, sys, threading, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vllm_metal_b200.metal import get_ops  # noqa: E402

SHAPES = {  # name: (bits, N, K)
    "8b qkv": (4, 6144, 4096), "8b o": (4, 4096, 4096), "8b gate_up": (4, 28672, 4096), "8b down": (4, 4096, 14336),
    "32b qkv": (8, 10240, 5120), "32b o": (8, 5120, 8192), "32b gate_up": (8, 51200, 5120), "32b down": (8, 5120, 25600),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ms", default="256,1024,2048,4096")
    ap.add_argument("--shapes", default="")
    a = ap.parse_args()
    ops = get_ops()
    peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"])
    clocks, stop = [], threading.Event()
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(0)

        def poll():
            while not stop.is_set():
                clocks.append((pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM),
                               pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)))
                time.sleep(0.005)
        threading.Thread(target=poll, daemon=True).start()
    except Exception:  # noqa: BLE001
        pass
    g = torch.Generator(device="cuda").manual_seed(0)
    print(f"# qgemm_tc, x fp16 -> y bf16, g64; burst bf16 peak {peak:.0f} TFLOP/s")
    for name, (bits, n, k) in SHAPES.items():
        if a.shapes and name not in a.shapes.split(","):
            continue
        copies = 3
        ws = []
        for _ in range(copies):
            w = torch.randint(-2**31, 2**31 - 1, (n, k * bits // 32), dtype=torch.int64, device="cuda", generator=g).to(torch.int32)
            s = (torch.rand((n, k // 64), device="cuda", generator=g) * 0.01 + 0.001).bfloat16()
            b = (-(2 ** (bits - 1)) * s.float()).bfloat16()
            ws.append((ops.qlinear_prepack(w, bits) if bits == 4 else w, s, b))
        line = f"{name:12s} [{n:6d} x {k:5d}] int{bits}:"
        for m in [int(t) for t in a.ms.split(",")]:
            x = torch.randn((m, k), device="cuda", generator=g).half()
            y = torch.empty((m, n), device="cuda", dtype=torch.bfloat16)

            def run():
                for wp, s, b in ws:
                    ops.quantized_matmul(x, wp, s, b, group_size=64, bits=bits, out=y, prepacked=True)
            run(); torch.cuda.synchronize()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                run()
            gr.replay(); torch.cuda.synchronize()
            n0 = len(clocks)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            reps = max(3, int(20000 / (m * n * k / 1e9 / 1.0 + 50)))
            for _ in range(reps):
                gr.replay()
            e1.record(); torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / (reps * copies)
            tf = 2.0 * m * n * k / us / 1e6
            mhz = sorted(c[0] for c in clocks[n0:]) or [0]
            line += f"  M={m}: {us:7.1f} us {tf:6.0f} TF/s ({tf / peak:.2f}, {mhz[len(mhz) // 2]} MHz)"
        print(line, flush=True)
    stop.set()
    reasons = set()
    for _, mask in clocks:
        for bit, nm in ((0x8, "hw_slowdown"), (0x40, "hw_thermal"), (0x20, "sw_thermal"), (0x4, "sw_power_cap")):
            if mask & bit:
                reasons.add(nm)
    print(f"# clock samples {len(clocks)}, throttle reasons seen: {sorted(reasons) or 'none'}")


if __name__ == "__main__":
    main()
