#!/usr/bin/env python
"""Headline benchmark: decode tok/s, Llama-3.1-8B affine int4 (g64), paged decode bs=64 ctx 4k.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

A "step" is one decode step of the whole model for the whole batch (32 layers of
quantized qkv -> RoPE -> reshape_and_cache -> paged attention -> o_proj -> SwiGLU MLP, final
norm, lm_head, greedy argmax) = 64 new tokens.  Synthetic data: random-init weights of the
real architecture, KV pool pre-filled with N(0,1) at ctx 4096, pages in a random permutation.

* ``value``      tokens/s with everything resident in HBM (CUDA-graph replay, device-side
                 token feedback), CUDA events, max over ranks.
* ``e2e``        same metric through the public runner API ``PagedModelRunner.decode_step``
                 with HOST inputs: per step host lists -> pinned -> H2D, graph, D2H token ids.
* ``roofline``   the dominant kernel (paged decode attention), timed live with CUDA events
                 over all 32 layers' caches (32 GiB working set >> L2), algorithmic bytes/launch.
* ``cpu_baseline`` the oracle's C port on the host cores over a bounded sample (rank 0, N=1).
* N > 1: tensor parallel over heads / MLP columns, NCCL allreduce after o_proj and down_proj
  (strong scaling: the global batch is fixed at 64).
* ``--impl reference``: the reference path's CPU restatement (oracle C port; the reference's
  own native code is Apple/Metal-only and cannot be built here) on the host cores.
"""

from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

MODEL = "llama-3.1-8b"
BATCH = 64
CTX = 4096
BLOCK = 16
METRIC = "decode tok/s (Llama-3.1-8B int4, bs=64)"
WORKLOAD = "Llama-3.1-8B affine int4 g64 random-init, paged decode bs=64 ctx 4096 block_size 16 (configs[1])"
MODEL_DESC = "Llama-3.1-8B (random-init, affine int4 g64 incl. lm_head)"

# --config: the other BASELINE.json configs as bench lines (the default, b, is the headline the driver runs).
# name: (preset, batch, ctx, metric, workload, model description, min gpus)
CONFIGS = {
    "b": (MODEL, BATCH, CTX, METRIC, WORKLOAD, MODEL_DESC, 1),
    "d": ("gpt-oss-20b-attn", 32, 4096, "decode tok/s (gpt-oss-20b attention blocks int4, bs=32)",
          "gpt-oss-20b attention blocks (sinks, 128-token window on even layers), int4 g64 q/k/v/o, paged decode "
          "bs=32 ctx 4096 block_size 16 (configs[3]; MoE MLP out of scope)",
          "gpt-oss-20b attention blocks (random-init, affine int4 g64 incl. lm_head)", 1),
    "c": ("qwen3-32b", 32, 8192, "mixed prefill+decode tok/s (Qwen3-32B int8, 2048-token chunk + 32 decode rows)",
          "Qwen3-32B affine int8 g64 random-init, mixed varlen batch: one 2048-token prefill chunk continuing at "
          "offset 6144 + 32 decode rows at ctx U[512, 8192], block_size 16, TP=2/4 (configs[2])",
          "Qwen3-32B (random-init, affine int8 g64 incl. lm_head)", 1),
    "e": ("llama-3.1-70b", 16, 32768, "decode tok/s (Llama-3.1-70B int4, bs=16, ctx 32k)",
          "Llama-3.1-70B affine int4 g64 random-init, split-KV paged decode bs=16 ctx 32768 block_size 16 at TP=8 "
          "(configs[4])", "Llama-3.1-70B (random-init, affine int4 g64 incl. lm_head)", 8),
}


def select_config(name: str) -> None:
    global MODEL, BATCH, CTX, METRIC, WORKLOAD, MODEL_DESC
    MODEL, BATCH, CTX, METRIC, WORKLOAD, MODEL_DESC, _ = CONFIGS[name]


def base_config(gpus: int) -> dict:
    """The workload description both arms print (key-identical dicts)."""
    return {"workload": WORKLOAD, "model": MODEL_DESC,
            "global_batch": BATCH, "ctx": CTX, "block_size": BLOCK, "kv_cache": "bf16",
            "parallelism": f"tp{gpus}", "decode_steps_timed": "steps",
            "l2_policy": "inputs larger than L2: every step streams the whole KV pool and all weights "
                         "(sizes in per_rank_resident)"}


def _ncu_traffic(summary: str) -> float | None:
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch from a committed
    ``ncu --set full`` summary under profiles/ (per launch, like ``achieved``); None if absent."""
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    total, seen = 0.0, 0
    try:
        with open(os.path.join(ROOT, "profiles", summary)) as f:
            for line in f:
                if line.startswith(("dram__bytes_read.sum", "dram__bytes_write.sum")):
                    unit = line.split("[")[1].split("]")[0]
                    total += float(line.split("=")[1]) * scale[unit]
                    seen += 1
    except Exception:  # noqa: BLE001
        return None
    return total if seen == 2 else None


def _peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback"


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the run (B200_PROFILING.md): an NVML polling
    thread (5 ms period) started before the warm-up; ``mark()``/``window()`` bracket the timed
    region on the host clock and only samples inside it are reported.  Falls back to an
    ``nvidia-smi -lms`` pipe if NVML cannot be loaded."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int) -> None:
        self.index, self.rows, self.proc, self._stop = index, [], None, threading.Event()
        self.mode, self.max_mhz, self.t0, self.t1 = "none", None, None, None

    def start(self):
        try:
            import pynvml

            pynvml.nvmlInit()
            # NVML enumerates physical devices: honour CUDA_VISIBLE_DEVICES when it is a plain index list
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            phys = self.index
            if vis and all(v.strip().isdigit() for v in vis.split(",")):
                phys = int(vis.split(",")[self.index])
            h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))

            def poll():
                while not self._stop.is_set():
                    try:
                        mhz = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
                        mask = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                        self.rows.append((time.perf_counter(), float(mhz), int(mask)))
                    except Exception:  # noqa: BLE001
                        pass
                    time.sleep(0.005)

            threading.Thread(target=poll, daemon=True).start()
            self.mode = "nvml"
            return
        except Exception:  # noqa: BLE001
            pass
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
            self.mode = "nvidia-smi"
        except Exception:  # noqa: BLE001
            self.proc = None

    def _pump(self):
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        bits = {v: k for k, v in self.REASONS.items()}
        for line in self.proc.stdout:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 6:
                continue
            try:
                mhz, self.max_mhz = float(f[0]), float(f[1])
            except ValueError:
                continue
            mask = 0
            for n, v in zip(names, f[2:6], strict=True):
                if v.lower().startswith("active"):
                    mask |= bits[n]
            self.rows.append((time.perf_counter(), mhz, mask))

    def mark(self):
        self.t0 = time.perf_counter()

    def window(self):
        self.t1 = time.perf_counter()

    def stop(self) -> dict:
        time.sleep(0.03)
        self._stop.set()
        if self.proc is not None:
            self.proc.terminate()
        if self.mode == "none":
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["clock sampler unavailable"], "samples": 0}
        inside = [r for r in self.rows if self.t0 is not None and self.t0 <= r[0] <= (self.t1 or 1e30)]
        rows = inside if inside else self.rows  # never empty-handed: say which set was used
        sm = sorted(r[1] for r in rows)
        reasons = set()
        for r in rows:
            for bit, name in self.REASONS.items():
                if r[2] & bit:
                    reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(reasons),
                "samples": len(inside), "samples_whole_run": len(self.rows), "source": self.mode,
                "scope": "timed region" if inside else "whole run (no sample fell inside the timed region)"}


# ------------------------------------------------------------------ CPU baseline ---

class CpuSample:
    """One decode layer (+ scaled lm_head) of the workload on the host cores, oracle C port."""

    def __init__(self, seed: int = 0, batch: int = BATCH, ctx: int = CTX) -> None:
        import numpy as np

        from oracle import cport
        from vllm_metal_b200.lm.decoder import PRESETS

        self.np, self.cport = np, cport
        # all host cores, whatever OMP_NUM_THREADS the launcher exported (torchrun sets it to 1)
        cport.set_num_threads(os.cpu_count() or 1)
        a = self.a = PRESETS[MODEL]
        rng = np.random.default_rng(seed)
        self.batch = batch
        hkv, hq, d = a.n_kv_heads, a.n_heads, a.head_dim
        nblk = batch * ((ctx + BLOCK) // BLOCK + 1)
        bf = lambda shape: rng.integers(0x3C00, 0x3F80, shape, dtype=np.uint16)  # noqa: E731  bf16 in [1, 2)
        self.bf = bf
        self.rng = rng
        self.kc, self.vc = bf((nblk, BLOCK, hkv, d)), bf((nblk, BLOCK, hkv, d))
        self.bt = rng.permutation(nblk).astype(np.int32).reshape(batch, -1)
        self.cu = np.arange(batch + 1, dtype=np.int32)
        self.seq = np.full(batch, ctx + 1, np.int32)
        self.slots = self.bt[np.arange(batch), ctx // BLOCK].astype(np.int64) * BLOCK + ctx % BLOCK
        self.pos = np.full(batch, ctx, np.int32)
        self.wqkv, self.wo = self.qw((hq + 2 * hkv) * d, a.hidden), self.qw(a.hidden, hq * d)
        self.wgu, self.wdn = self.qw(2 * a.intermediate, a.hidden), self.qw(a.hidden, a.intermediate)
        self.norm_w = bf((a.hidden,))
        self.h = bf((batch, a.hidden))
        self.threads = cport.num_threads()

    def qw(self, n, k):
        np, a = self.np, self.a
        return (self.rng.integers(0, 2**32, (n, k // 8), dtype=np.uint32), self.bf((n, k // a.group_size)),
                self.bf((n, k // a.group_size)))

    def layer(self) -> float:
        np, cport, a, batch = self.np, self.cport, self.a, self.batch
        hkv, hq, d, g, b = a.n_kv_heads, a.n_heads, a.head_dim, a.group_size, a.bits
        t0 = time.perf_counter()
        x = cport.rms_norm(self.h, self.norm_w, a.rms_eps)
        qkv = cport.qmatmul(x, *self.wqkv, g, b)
        q = np.ascontiguousarray(qkv[:, :hq * d]).reshape(batch, hq, d)
        k = np.ascontiguousarray(qkv[:, hq * d:(hq + hkv) * d]).reshape(batch, hkv, d)
        v = np.ascontiguousarray(qkv[:, (hq + hkv) * d:]).reshape(batch, hkv, d)
        cport.rope_(q, self.pos, a.rope_base)
        cport.rope_(k, self.pos, a.rope_base)
        cport.reshape_and_cache(k, v, self.kc, self.vc, self.slots)
        o = cport.paged_attention(q, self.kc, self.vc, self.bt, self.seq, self.cu, d ** -0.5)
        att = cport.qmatmul(o.reshape(batch, hq * d), *self.wo, g, b)
        res = self.h.copy()
        x = cport.rms_norm(att, self.norm_w, a.rms_eps, residual=res)
        cport.qmatmul(cport.silu_mul(cport.qmatmul(x, *self.wgu, g, b)), *self.wdn, g, b)
        return time.perf_counter() - t0

    def head(self) -> float:
        """lm_head + argmax on 1/8 of the vocab rows, scaled to the full vocab."""
        a = self.a
        rows = a.vocab // 8
        whead = self.qw(rows, a.hidden)
        t0 = time.perf_counter()
        logits = self.cport.qmatmul(self.h, *whead, a.group_size, a.bits)
        self.cport.argmax(logits)
        return (time.perf_counter() - t0) * (a.vocab / rows)


def run_reference(args) -> None:
    """--impl reference: the path's CPU restatement on the host cores, rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    cs = CpuSample()
    a = cs.a
    for _ in range(args.warmup):
        cs.layer()
    t_head = cs.head()
    # each "step" = a bounded sample of one decode step: 1 of 32 layers
    times = [cs.layer() for _ in range(args.steps)]
    t_layer = sum(times) / len(times)
    step_s = a.num_layers * t_layer + t_head
    val = BATCH / step_s
    sample = (f"each step times 1 of {a.num_layers} layers (attention over 64x{CTX}-token bf16 KV + qkv/o/gate_up/down "
              f"int4 g64 linears at M=64) with the oracle C port; step = {a.num_layers} x layer + lm_head "
              f"(timed once on 1/8 of the vocab rows, x8)")
    _emit(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": "tok/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": step_s * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": base_config(args.gpus),
        "note": "reference native code is Apple/Metal-only (unbuildable here); this arm is the CPU restatement of "
                "the same path (oracle C port, OpenMP, all host cores on rank 0)",
        "cpu_baseline": {"value": val, "unit": "tok/s", "cores": cs.threads, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# ------------------------------------------------------------------------ GPU arm ---

def run_mixed(args) -> None:
    """Config c: a mixed prefill+decode step (eager launches through ``PagedModelRunner.execute``: host lists in,
    token ids out every step, so ``value`` already IS the end-to-end number and ``e2e`` repeats the same run's
    bytes).  M > 64 linears run the tensor-core GEMM (csrc/qgemm_tc.cu), the chunk's attention the tcgen05 tile
    kernel, the decode rows the per-row kernels."""
    import numpy as np
    import torch
    import torch.distributed as dist

    from vllm_metal_b200.attention.caches.kv_cache import B200PagedKVCache
    from vllm_metal_b200.lm.decoder import PRESETS, Model
    from vllm_metal_b200.metal import get_ops
    from vllm_metal_b200.v1.model_runner import PagedModelRunner

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    ops = get_ops()
    a = PRESETS[MODEL]
    chunk, offset, ndec = 2048, int(os.environ.get("B200_CHUNK_OFFSET", "6144")), BATCH
    T = chunk + ndec
    nvl = None
    if world > 1:
        from vllm_metal_b200.distributed.nvlink_allreduce import NVLinkAllReduce

        nvl = NVLinkAllReduce(device=dev, max_bytes=T * a.hidden * 2)

    model = Model(a, torch.bfloat16, dev, seed=0, rank=rank, world=world, allreduce=nvl, prefill_layout=True)
    rng = np.random.default_rng(4)
    dec_ctx = rng.integers(512, CTX, ndec).tolist()
    lens = [n + 1 for n in dec_ctx] + [offset + chunk]
    per = [(n + BLOCK - 1) // BLOCK for n in lens]
    nb = 1 + sum(per)
    cache = B200PagedKVCache(a.num_layers, a.n_kv_heads // world, a.head_dim, nb, BLOCK, torch.bfloat16, dev)
    g = torch.Generator(device=dev).manual_seed(3 + rank)
    for t in (cache._k_all, cache._v_all):
        for i in range(a.num_layers):
            t[i].normal_(0.0, 1.0, generator=g)
    runner = PagedModelRunner(model, cache, max_batch=ndec + 1, max_model_len=max(lens) + BLOCK, use_cuda_graph=False)
    perm = (rng.permutation(nb - 1) + 1).tolist()
    tables, o = [], 0
    for n in per:
        tables.append(perm[o:o + n])
        o += n
    dec = [(tables[i], dec_ctx[i]) for i in range(ndec)]
    pre = [(tables[ndec], chunk, offset)]
    ids = rng.integers(0, a.vocab, T).tolist()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local)
    sampler.start()
    for _ in range(args.warmup):
        runner.execute(ids, dec, pre)
    l0 = ops.launches
    h2d0, d2h0 = runner.h2d_bytes, runner.d2h_bytes
    barrier()
    sampler.mark()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        runner.execute(ids, dec, pre)
    e1.record()
    barrier()
    sampler.window()
    clocks = sampler.stop()
    ms = e0.elapsed_time(e1)
    launches = ops.launches - l0
    h2d, d2h = (runner.h2d_bytes - h2d0) // args.steps, (runner.d2h_bytes - d2h0) // args.steps

    # dominant kernels, timed alone: the gate_up GEMM at M = T (tensor bound) and the chunk's attention
    gu = model.layers[0].mlp.gate_up_proj
    x = torch.randn((T, a.hidden), device=dev, generator=g).bfloat16()
    for _ in range(2):
        for layer in model.layers[:8]:
            layer.mlp.gate_up_proj(x)
    barrier()
    e0.record()
    for layer in model.layers[:16]:
        layer.mlp.gate_up_proj(x)
    e1.record()
    barrier()
    gemm_ms = e0.elapsed_time(e1) / 16
    gemm_flops = 2.0 * T * gu.out_features * gu.in_features
    hq, hkv, d = a.n_heads // world, a.n_kv_heads // world, a.head_dim
    q = torch.randn((chunk, hq, d), device=dev, generator=g).bfloat16()
    out = torch.empty_like(q)
    bt = torch.tensor([tables[ndec]], dtype=torch.int32, device=dev)
    cu = torch.tensor([0, chunk], dtype=torch.int32, device=dev)
    sl = torch.tensor([offset + chunk], dtype=torch.int32, device=dev)
    for _ in range(3):
        ops.paged_attention_primitive(q, cache.key_caches[0], cache.value_caches[0], hkv, d ** -0.5, 0.0, bt, sl, cu, BLOCK,
                                      offset + chunk, -1, out)
    barrier()
    e0.record()
    for i in range(16):
        ops.paged_attention_primitive(q, cache.key_caches[i], cache.value_caches[i], hkv, d ** -0.5, 0.0, bt, sl, cu,
                                      BLOCK, offset + chunk, -1, out)
    e1.record()
    barrier()
    attn_ms = e0.elapsed_time(e1) / 16
    attn_flops = 4.0 * d * hq * chunk * (offset + (chunk + 1) / 2)
    times = torch.tensor([ms, gemm_ms, attn_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    ms, gemm_ms, attn_ms = times.tolist()
    if rank == 0:
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                pk = json.load(f)
            peak_tf, peak_src = float(pk.get("bf16_tflops_sustained", pk["bf16_tflops"])), "measured (sustained)"
        except Exception:  # noqa: BLE001
            peak_tf, peak_src = 1400.0, "fallback"
        tok_s = T * args.steps / (ms * 1e-3)
        cfg = base_config(world)
        cfg.update({"global_batch": f"{ndec} decode rows + 1 chunk of {chunk}", "chunk_offset": offset,
                    "tokens_per_step": T, "ctx": f"decode U[512,{CTX}] (seed 4), chunk ctx {offset + chunk}"})
        _emit(json.dumps({
            "metric": METRIC, "value": tok_s, "unit": "tok/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": cfg,
            "allreduce": "none" if world == 1 else f"nvlink kernel, {nvl.calls} calls, {nvl.nccl_calls} routed to NCCL",
            "per_rank_resident": f"KV pool {cache.nbytes / 2**30:.1f} GiB + weights {model.weight_bytes() / 2**30:.2f} GiB "
                                 f"(stream-K blobs + tensor-core GEMM layout)",
            "e2e": {"value": tok_s, "unit": "tok/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "note": "the timed loop is the public runner API with host lists in and token ids out"},
            "gpu_launches": int(launches), "clocks": clocks,
            "roofline": {"bound": "tensor", "kernel": f"qgemm_tc_kernel gate_up {gu.out_features}x{gu.in_features} int{a.bits} "
                                                      f"g{a.group_size}, M={T} (+ cast + silu_mul)",
                         "achieved": gemm_flops / (gemm_ms * 1e-3) / 1e12, "peak": peak_tf, "peak_source": peak_src,
                         "unit": "TFLOP/s", "frac": gemm_flops / (gemm_ms * 1e-3) / 1e12 / peak_tf, "traffic": None,
                         "flops_per_launch": gemm_flops, "us_per_launch": gemm_ms * 1e3},
            "roofline_attention": {"bound": "tensor", "kernel": "prefill_tc_kernel (2048-token chunk, causal)",
                                   "achieved": attn_flops / (attn_ms * 1e-3) / 1e12, "peak": peak_tf, "unit": "TFLOP/s",
                                   "frac": attn_flops / (attn_ms * 1e-3) / 1e12 / peak_tf, "flops_per_launch": attn_flops,
                                   "us_per_launch": attn_ms * 1e3},
            "cpu_baseline": None,
        }))
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


def run_b200(args) -> None:
    import numpy as np
    import torch
    import torch.distributed as dist

    from vllm_metal_b200.attention.caches.kv_cache import B200PagedKVCache
    from vllm_metal_b200.lm.decoder import PRESETS, Model
    from vllm_metal_b200.metal import get_ops
    from vllm_metal_b200.v1.model_runner import PagedModelRunner

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert world == args.gpus or world == 1, f"WORLD_SIZE={world} but --gpus {args.gpus}"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    ops = get_ops()
    a = PRESETS[MODEL]
    nvl = None
    if world > 1 and os.environ.get("B200_ALLREDUCE", "nvlink") != "nccl":
        from vllm_metal_b200.distributed.nvlink_allreduce import NVLinkAllReduce

        nvl = NVLinkAllReduce(device=dev, max_bytes=BATCH * a.hidden * 2)

    def nccl_allreduce(x):
        dist.all_reduce(x)

    # the NVLink object itself: the decoder then uses its fused all-reduce + residual + RMSNorm kernel
    model = Model(a, torch.bfloat16, dev, seed=0, rank=rank, world=world,
                  allreduce=nvl if nvl is not None else nccl_allreduce)
    total = args.warmup + args.steps
    max_len = CTX + 2 * total + 8
    blocks_per_seq = (max_len + BLOCK - 1) // BLOCK
    nb = 1 + BATCH * blocks_per_seq
    cache = B200PagedKVCache(a.num_layers, a.n_kv_heads // world, a.head_dim, nb, BLOCK, torch.bfloat16, dev,
                             sliding_window_per_layer=a.sliding_window_per_layer())
    g = torch.Generator(device=dev).manual_seed(3 + rank)
    for t in (cache._k_all, cache._v_all):
        for i in range(a.num_layers):
            t[i].normal_(0.0, 1.0, generator=g)
    runner = PagedModelRunner(model, cache, max_batch=BATCH, max_model_len=blocks_per_seq * BLOCK,
                              use_cuda_graph=True)
    rng = np.random.default_rng(2)
    tables = (rng.permutation(nb - 1) + 1).astype(np.int32).reshape(BATCH, blocks_per_seq)
    tokens = rng.integers(0, a.vocab, BATCH).tolist()
    seq_lens = [CTX] * BATCH

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- e2e through the public runner API (host inputs every step) ----------
    for _ in range(args.warmup):
        tokens = runner.decode_step(tokens, tables, seq_lens)
        seq_lens = [n + 1 for n in seq_lens]
    h2d0, d2h0 = runner.h2d_bytes, runner.d2h_bytes
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        tokens = runner.decode_step(tokens, tables, seq_lens)
        seq_lens = [n + 1 for n in seq_lens]
    e1.record()
    barrier()
    e2e_ms = e0.elapsed_time(e1)
    h2d = (runner.h2d_bytes - h2d0) // args.steps
    d2h = (runner.d2h_bytes - d2h0) // args.steps

    # ---------------- device-resident loop: graph replay + on-device feedback --------------
    max_len_now = runner.fill_decode(tokens, tables, seq_lens)
    runner.buf.upload()
    bucket_len = max_len_now + total  # same split-KV bucket for the whole run
    sampler = ClockSampler(local)
    sampler.start()  # before the warm-up, so the first samples exist when the timed region opens
    for _ in range(args.warmup):
        runner.run_decode(BATCH, bucket_len, advance=True)
    barrier()
    sampler.mark()
    e0.record()
    for _ in range(args.steps):
        runner.run_decode(BATCH, bucket_len, advance=True)
    e1.record()
    barrier()
    sampler.window()
    clocks = sampler.stop()
    dev_ms = e0.elapsed_time(e1)
    if os.environ.get("B200_PROFILE_RANGE"):
        # one extra, untimed step inside a cudaProfilerStart/Stop range for
        # `ncu --profile-from-start off` (profiles/: launch list + --set full capture)
        torch.cuda.profiler.start()
        runner.run_decode(BATCH, bucket_len, advance=True)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
    per_step_launches = runner.launches_per_graph.get((BATCH, runner._bucket(bucket_len), True), 0)

    # ---------------- dominant kernel: paged decode attention, all layers' caches ----------
    hq, hkv, d = a.n_heads // world, a.n_kv_heads // world, a.head_dim
    meta = runner._meta(BATCH, runner._bucket(bucket_len))
    q = torch.randn((BATCH, hq, d), device=dev, generator=g).bfloat16()
    out = torch.empty_like(q)
    cur_len = int(runner.buf.view("seq_lens")[:BATCH].max().item())

    def attn_all_layers():
        for i in range(a.num_layers):
            ops.paged_attention_primitive(q, cache.key_caches[i], cache.value_caches[i], hkv, d ** -0.5, 0.0,
                                          meta.block_tables, meta.seq_lens, meta.cu_seqlens_q, BLOCK,
                                          meta.max_seq_len, -1, out)

    def timed_graph(fn, reps):
        """fn() captured once in a CUDA graph and replayed: the host's launch cost (~20 us per call through
        ctypes) must not pace kernels that take less than that (it did at TP >= 2 in round 1)."""
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            fn()
        gr.replay()
        barrier()
        e0.record()
        for _ in range(reps):
            gr.replay()
        e1.record()
        barrier()
        return e0.elapsed_time(e1) / reps

    reps = max(3, min(args.steps, 10))
    attn_ms = timed_graph(attn_all_layers, reps) / a.num_layers
    alg_bytes = BATCH * (2 * cur_len * hkv * d * 2 + 2 * hq * d * 2 + ((cur_len + BLOCK - 1) // BLOCK) * 4)
    peak, peak_src = _peaks()
    achieved = alg_bytes / (attn_ms * 1e-3) / 1e9

    # ---------------- secondary: int4 GEMV (gate_up + fused SwiGLU, stream-K kernel) --------
    # every layer's own blob in turn (32 x 57 MB >> L2), x in fp16 as the decoder's norm hands it over
    gemv_ms, gemv_bytes, gemv_gbs = 0.0, 0.0, 0.0
    mlp = model.layers[0].mlp
    if mlp is not None:
        x = torch.randn((BATCH, a.hidden), device=dev, generator=g).half()
        def gate_up_all_layers():
            for layer in model.layers:
                layer.mlp.gate_up_proj(x, out_f16=True)

        gemv_ms = timed_graph(gate_up_all_layers, reps) / a.num_layers
        n_gu, k_gu = mlp.gate_up_proj.out_features, mlp.gate_up_proj.in_features
        n_out = n_gu // 2 if mlp.gate_up_proj.swiglu else n_gu
        gemv_bytes = n_gu * k_gu * (a.bits / 8 + 4 / a.group_size) + BATCH * (k_gu + n_out) * 2
        gemv_gbs = gemv_bytes / (gemv_ms * 1e-3) / 1e9

    # max over ranks
    times = torch.tensor([dev_ms, e2e_ms, attn_ms, gemv_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms, attn_ms, gemv_ms = times.tolist()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline and args.config == "b":
        layers = 6
        cs = CpuSample()
        cs.layer()
        t_layer = sum(cs.layer() for _ in range(layers)) / layers
        t_head, threads = cs.head(), cs.threads
        step_s = a.num_layers * t_layer + t_head
        cpu = {"value": BATCH / step_s, "unit": "tok/s", "cores": threads, "kind": "port",
               "sample": f"{layers} of {a.num_layers} decode layers (bs=64, ctx 4096: int4 g64 linears + paged "
                         f"attention over bf16 KV) and the lm_head on 1/8 of the vocab, timed with the oracle C "
                         f"port (gcc -O3 -fopenmp), scaled to {a.num_layers} layers + full vocab"}

    if rank == 0:
        tok_s = BATCH * args.steps / (dev_ms * 1e-3)
        line = {
            "metric": METRIC, "value": tok_s, "unit": "tok/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": base_config(world),
            "allreduce": ("none" if world == 1 else "nvlink kernel fused with residual add + RMSNorm (csrc/nvlink_allreduce.cu)"
                          if nvl is not None else "nccl"),
            "per_rank_resident": f"KV pool {cache.nbytes / 2**30:.1f} GiB + weights {model.weight_bytes() / 2**30:.2f} GiB",
            "e2e": {"value": BATCH * args.steps / (e2e_ms * 1e-3), "unit": "tok/s",
                    "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)},
            "gpu_launches": int(per_step_launches * args.steps),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": "paged_decode_mma_kernel (+ split_reduce_kernel when the plan splits)",
                         "achieved": achieved, "peak": peak, "peak_source": peak_src, "unit": "GB/s",
                         "frac": achieved / peak, "frac_of_nominal_8000": achieved / 8000.0,
                         "traffic": _ncu_traffic("r02_decode_attn_full_summary.txt") if world == 1 and args.config == "b" else None,
                         "traffic_source": "profiles/r02_decode_attn_full_summary.txt (ncu --set full of the shipped "
                                           "kernel at this shape, single-pass plan; dram read + write per launch)",
                         "bytes_per_launch": alg_bytes, "us_per_launch": attn_ms * 1e3},
            "roofline_qgemv": None if mlp is None else {
                "bound": "hbm",
                "kernel": f"qgemm_stream_kernel, persistent stream-K, fp16 A in TMEM, SwiGLU epilogue "
                          f"(gate_up {n_gu}x{k_gu} int{a.bits} g{a.group_size}, M={BATCH})",
                "achieved": gemv_gbs, "peak": peak, "unit": "GB/s", "frac": gemv_gbs / peak,
                "traffic": _ncu_traffic("r02_qstream_gate_up_full_summary.txt") if world == 1 and args.config == "b" else None,
                "traffic_source": "profiles/r02_qstream_gate_up_full_summary.txt (ncu --set full, same shape)",
                "bytes_per_launch": gemv_bytes, "us_per_launch": gemv_ms * 1e3},
            "cpu_baseline": cpu,
        }
        _emit(json.dumps(line))
    if world > 1:
        # NCCL collectives live inside captured CUDA graphs: tearing the communicator down with
        # those graphs alive can block forever (seen on 2xB200), so every rank syncs and leaves
        # without running the destructors.
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


_REAL_STDOUT = None


def _capture_stdout() -> None:
    """Everything libraries print to fd 1 (NCCL's version banner, ...) goes to stderr; the one JSON
    line is written to the real stdout by ``_emit``."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def _emit(text: str) -> None:
    sys.stdout.flush()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, (text + "\n").encode())


def main() -> None:
    _capture_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=128)  # >= 1 s timed at the headline config
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--config", default="b", choices=sorted(CONFIGS), help="BASELINE.json config (default b = headline)")
    args = ap.parse_args()
    select_config(args.config)
    if args.gpus < CONFIGS[args.config][6] and args.impl == "b200":
        _emit(json.dumps({"config": args.config, "unavailable": f"config {args.config} needs --gpus >= {CONFIGS[args.config][6]}"}))
        return
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    elif args.config == "c":
        if args.steps == 128:
            args.steps = 8  # a mixed step is ~2080 tokens; 8 steps already time >= 1 s of work per GPU
        run_mixed(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
