#!/usr/bin/env python
"""Generate golden fixtures by running the REFERENCE's own Python code.

Run in the build container only (``/root/reference`` does not exist on the GPU
box):  ``python tests/golden/make_golden.py``.  Writes

* ``tests/golden/metadata_golden.json`` -- outputs of the reference's
  ``prepare_grouped`` / ``prepare_unified`` (attention/context.py:162-314) and
  ``_pick_kernel_block_size`` / ``_build_block_tables``
  (attention/impls/sdpa.py:115-163) on seeded random request batches.
* ``tests/golden/attention_golden.npz`` -- outputs of the reference's test oracle
  ``ref_paged_attn`` (tools/attention_bench_utils.py:11-62) on seeded inputs.

MLX is not installable here, so ``mlx.core`` is replaced by a NumPy stand-in
exposing only the array ops those functions call (array/zeros/ones/arange/
einsum/triu/where/softmax/tanh/repeat/concatenate/logical_*); the control flow,
indexing, masks and scaling executed are the reference's own source files, read
from /root/reference at generation time (nothing is copied into this repo).
"""

from __future__ import annotations

import importlib.util
import json
import os
import random
import sys
import types

import numpy as np

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def _install_stubs() -> None:
    mx = types.ModuleType("mlx.core")
    mx.array = lambda v, dtype=None: np.array(v, dtype=dtype)
    mx.zeros = lambda shape, dtype=None: np.zeros(shape, dtype=dtype)
    mx.ones = lambda shape, dtype=np.float32: np.ones(shape, dtype=dtype)
    mx.arange = lambda *a, dtype=None: np.arange(*a, dtype=dtype)
    mx.int32, mx.int64, mx.float32, mx.float16, mx.bool_ = (
        np.int32,
        np.int64,
        np.float32,
        np.float16,
        np.bool_,
    )
    mx.uint32 = np.uint32
    mx.einsum = lambda spec, *ops: np.einsum(spec, *ops)
    mx.triu = lambda a, k=0: np.triu(a, k=k)
    mx.logical_not = np.logical_not
    mx.logical_or = np.logical_or
    mx.tanh = np.tanh
    mx.where = np.where
    mx.repeat = lambda a, n, axis=None: np.repeat(a, n, axis=axis)
    mx.concatenate = lambda xs, axis=0: np.concatenate(xs, axis=axis)

    def softmax(a, axis=-1):
        m = a.max(axis=axis, keepdims=True)
        m = np.where(np.isfinite(m), m, 0.0)
        e = np.exp(a - m)
        return e / e.sum(axis=axis, keepdims=True)

    mx.softmax = softmax
    mx.Dtype = type
    mlx = types.ModuleType("mlx")
    mlx.core = mx
    nn = types.ModuleType("mlx.nn")
    nn.Module = object
    mlx.nn = nn
    sys.modules.update({"mlx": mlx, "mlx.core": mx, "mlx.nn": nn})

    for name in ("mlx_lm", "mlx_lm.models", "mlx_lm.models.base"):
        sys.modules[name] = types.ModuleType(name)
    sys.modules["mlx_lm.models.base"].create_causal_mask = lambda *a, **k: None

    # package shells so the reference modules can be loaded by path without
    # running vllm_metal/__init__.py (which needs vLLM + MLX at import time)
    for name in (
        "vllm_metal",
        "vllm_metal.attention",
        "vllm_metal.attention.caches",
        "vllm_metal.attention.impls",
    ):
        m = types.ModuleType(name)
        m.__path__ = []  # mark as package
        sys.modules[name] = m
    kvc = types.ModuleType("vllm_metal.attention.caches.kv_cache")
    kvc.MetalPagedKVCache = object
    sys.modules[kvc.__name__] = kvc
    rope = types.ModuleType("vllm_metal.attention.impls.varlen_rope_compat")
    rope.apply_attention_rope = None
    sys.modules[rope.__name__] = rope
    metal = types.ModuleType("vllm_metal.metal")
    metal.get_ops = None
    sys.modules[metal.__name__] = metal


def _load(name: str, path: str):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def _rand_requests(rng: random.Random, block_sizes):
    """Random decode + prefill requests with per-group block ids."""
    dec, pre = [], []
    for _ in range(rng.randint(0, 4)):
        seq_len = rng.randint(0, 70)
        n = rng.choice([1, 1, 1, 2, 3])
        ids = [
            [rng.randint(0, 99) for _ in range((seq_len + n + bs - 1) // bs + rng.randint(0, 1))]
            for bs in block_sizes
        ]
        dec.append((ids, seq_len) if n == 1 and rng.random() < 0.5 else (ids, seq_len, n))
    for _ in range(rng.randint(0, 3)):
        start = rng.choice([0, 0, rng.randint(1, 40)])
        n = rng.randint(1, 40)
        ids = [
            [rng.randint(0, 99) for _ in range((start + n + bs - 1) // bs)] for bs in block_sizes
        ]
        pre.append((ids, n, start))
    if not dec and not pre:
        pre.append(([[rng.randint(0, 99)] for _ in block_sizes], 1, 0))
    return dec, pre


def make_metadata(context, sdpa) -> dict:
    rng = random.Random(20261104)
    cases = []
    for ci in range(60):
        block_sizes = tuple(rng.choice([[4], [8], [16], [32], [16, 32], [32, 16], [16, 16]]))
        dec, pre = _rand_requests(rng, block_sizes)
        merge = rng.random() < 0.3
        context.prepare_grouped(dec, pre, block_sizes, merge_verify_windows=merge)
        ctx = context.get_context()
        cases.append(
            {
                "decode": dec,
                "prefill": pre,
                "block_sizes": list(block_sizes),
                "merge": merge,
                "out": {
                    "slot_mapping": ctx.slot_mapping,
                    "block_tables": ctx.block_tables,
                    "context_lens": ctx.context_lens,
                    "cu_seqlens": ctx.cu_seqlens,
                    "offsets": ctx.offsets,
                    "num_decode_requests": ctx.num_decode_requests,
                    "verify_window_q": ctx.verify_window_q,
                    "kv_groups": [
                        {
                            "slot_mapping": g.slot_mapping,
                            "block_tables": g.block_tables,
                            "block_size": g.block_size,
                        }
                        for g in ctx.kv_groups
                    ],
                },
            }
        )
        context.clear_context()
    tables = []
    for cbs in (8, 16, 32, 24, 48, 64, 96, 128, 544):
        for _ in range(3):
            raw = [
                [rng.randint(0, 50) for _ in range(rng.randint(1, 5))]
                for _ in range(rng.randint(1, 4))
            ]
            bt, kbs = sdpa._build_block_tables(raw, cbs)
            tables.append(
                {"raw": raw, "cache_block_size": cbs, "tables": np.asarray(bt).tolist(), "kernel_block_size": kbs}
            )
    picks = []
    for cbs in list(range(1, 80)) + [128, 256, 544, 1000]:
        try:
            picks.append([cbs, sdpa._pick_kernel_block_size(cbs)])
        except ValueError:
            picks.append([cbs, None])
    return {"prepare": cases, "build_block_tables": tables, "pick_kernel_block_size": picks}


def make_attention(bench) -> dict[str, np.ndarray]:
    rng = np.random.default_rng(7)
    out: dict[str, np.ndarray] = {}
    specs = [
        # name, seq (q_len, kv_len) list, Hq, Hkv, D, block, window, softcap, dtype
        ("decode_gqa", [(1, 37), (1, 1), (1, 130)], 8, 2, 64, 16, None, None, np.float32),
        ("decode_window", [(1, 300), (1, 20)], 4, 4, 64, 16, 128, None, np.float32),
        ("mixed", [(1, 99), (5, 18), (33, 70)], 8, 2, 64, 16, None, None, np.float32),
        ("mixed_window", [(3, 150), (17, 140)], 4, 2, 32, 8, 24, None, np.float32),
        ("softcap", [(4, 40)], 4, 1, 32, 32, None, 30.0, np.float32),
        ("half", [(1, 77), (9, 64)], 8, 4, 64, 16, None, None, np.float16),
    ]
    for name, seqs, hq, hkv, d, bs, win, cap, dt in specs:
        q_lens = [s[0] for s in seqs]
        kv_lens = [s[1] for s in seqs]
        nblk = 12
        kc = rng.standard_normal((nblk, bs, hkv, d)).astype(dt)
        vc = rng.standard_normal((nblk, bs, hkv, d)).astype(dt)
        q = rng.standard_normal((sum(q_lens), hq, d)).astype(dt)
        mb = (max(kv_lens) + bs - 1) // bs
        bt = rng.integers(0, nblk, size=(len(seqs), mb)).astype(np.int32)
        scale = float(d) ** -0.5
        ref = bench.ref_paged_attn(
            q, kc, vc, q_lens, kv_lens, bt, scale, sliding_window=win, soft_cap=cap
        )
        out[f"{name}.q"] = q
        out[f"{name}.kc"] = kc
        out[f"{name}.vc"] = vc
        out[f"{name}.bt"] = bt
        out[f"{name}.meta"] = np.array(
            [hq, hkv, d, bs, -1 if win is None else win], dtype=np.int64
        )
        out[f"{name}.softcap"] = np.array([0.0 if cap is None else cap], dtype=np.float64)
        out[f"{name}.q_lens"] = np.array(q_lens, dtype=np.int64)
        out[f"{name}.kv_lens"] = np.array(kv_lens, dtype=np.int64)
        out[f"{name}.ref"] = np.asarray(ref)
    return out


def main() -> None:
    _install_stubs()
    context = _load("vllm_metal.attention.context", f"{REF}/vllm_metal/attention/context.py")
    sdpa = _load("vllm_metal.attention.impls.sdpa", f"{REF}/vllm_metal/attention/impls/sdpa.py")
    bench = _load("ref_attention_bench_utils", f"{REF}/tools/attention_bench_utils.py")
    with open(os.path.join(HERE, "metadata_golden.json"), "w") as f:
        json.dump(make_metadata(context, sdpa), f)
    np.savez_compressed(os.path.join(HERE, "attention_golden.npz"), **make_attention(bench))
    print("wrote metadata_golden.json, attention_golden.npz")


if __name__ == "__main__":
    main()
