"""Multi-GPU check of the NVLink one-shot all-reduce (run under torchrun on N = 2/4/8 B200s):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29533 tests/multi_gpu_allreduce.py

Not collected by pytest (needs several GPUs); the driver's single-GPU ``-m gpu`` tier cannot run
it.  Checks, for several sizes and dtypes: equal to an fp32 sum of the gathered inputs rounded once
(the kernel's contract), within the summation-order error bound of NCCL's result, bit-identical across ranks,
identical when replayed from a CUDA graph; then times it against ``dist.all_reduce``.
"""

from __future__ import annotations

import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main() -> None:
    rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from vllm_metal_b200.distributed.nvlink_allreduce import NVLinkAllReduce

    ar = NVLinkAllReduce(device=dev, max_bytes=1 << 20)
    ok = True
    for dtype in (torch.bfloat16, torch.float16, torch.float32):
        for numel in (8, 4096, 64 * 4096, 64 * 4096 + 8, (1 << 20) // 4):
            if numel * torch.empty((), dtype=dtype).element_size() > (1 << 20):
                continue
            g = torch.Generator(device=dev).manual_seed(1000 * rank + numel)
            x = torch.randn((numel,), device=dev, generator=g).to(dtype)
            gathered = [torch.empty_like(x) for _ in range(world)]
            dist.all_gather(gathered, x)
            want = torch.zeros((numel,), device=dev, dtype=torch.float32)
            for t in gathered:  # rank order, fp32, one rounding: the kernel's contract
                want += t.float()
            want = want.to(dtype)
            y = x.clone()
            for _ in range(3):  # repeated calls exercise both staging halves and the epoch counters
                y = x.clone()
                ar(y)
            torch.cuda.synchronize()
            if not torch.equal(y, want):
                ok = False
                print(f"[rank {rank}] MISMATCH dtype={dtype} numel={numel} max diff "
                      f"{(y.float() - want.float()).abs().max().item()}", flush=True)
            ref = x.clone()
            dist.all_reduce(ref)
            # NCCL sums in the storage dtype in its own (tree/ring) order: any order is within
            # (world) roundings of the running partial, each <= eps * sum_r |x_r|
            eps = {torch.float32: 2.0 ** -23, torch.float16: 2.0 ** -10, torch.bfloat16: 2.0 ** -7}[dtype]
            bound = world * eps * sum(t.float().abs() for t in gathered) + 1e-6
            if not bool(((y.float() - ref.float()).abs() <= bound).all()):
                ok = False
                print(f"[rank {rank}] differs from NCCL dtype={dtype} numel={numel}", flush=True)
            ys = [torch.empty_like(y) for _ in range(world)]
            dist.all_gather(ys, y)
            if not all(torch.equal(ys[0], t) for t in ys):
                ok = False
                print(f"[rank {rank}] ranks disagree dtype={dtype} numel={numel}", flush=True)
    assert ar.calls > 0 and ar.nccl_calls == 0, (ar.calls, ar.nccl_calls)

    # CUDA graph replay + timing at the headline message (64 x 4096 bf16 = 512 KiB)
    x0 = torch.randn((64, 4096), device=dev).bfloat16()
    bufs = [x0.clone() for _ in range(16)]

    def timed(fn, label):
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            for b in bufs:
                fn(b)
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=s):
                for b in bufs:
                    fn(b)
            graph.replay()
            torch.cuda.synchronize()
            dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(20):
                graph.replay()
            e1.record()
            torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / (20 * len(bufs))
        t = torch.tensor([us], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            print(f"{label}: {t.item():.2f} us per call (64 x 4096 bf16 = 512 KiB unless noted; graph replay, max over {world} ranks)", flush=True)

    for b in bufs:
        b.copy_(x0)
    timed(lambda t: ar(t), "nvlink one-shot")
    timed(lambda t: dist.all_reduce(t), "nccl")

    # ---- fused all-reduce + residual add + RMSNorm vs the two-step sequence (bit-identical at hidden 4096 / 8192,
    # where both kernels map threads to elements the same way; equal up to the fp32 order of the mean square else)
    from vllm_metal_b200.metal import get_ops

    ops = get_ops()
    for rows, hidden, f16 in ((64, 4096, True), (64, 4096, False), (1, 4096, True), (16, 8192, True), (37, 5120, False),
                              (100, 2880, True)):
        g = torch.Generator(device=dev).manual_seed(7 * rank + rows + hidden)
        part = torch.randn((rows, hidden), device=dev, generator=g).bfloat16()
        g2_ = torch.Generator(device=dev).manual_seed(rows + hidden)  # residual / weight: same on every rank
        res0 = torch.randn((rows, hidden), device=dev, generator=g2_).bfloat16()
        w = (torch.rand((hidden,), device=dev, generator=g2_) + 0.5).bfloat16()
        odt = torch.float16 if f16 else None
        # two-step reference
        s_ = part.clone()
        ar(s_)
        res_ref = res0.clone()
        want = ops.rms_norm(s_, w, 1e-5, residual=res_ref, out_dtype=odt if hidden >= 512 else None)
        for _ in range(2):
            res = res0.clone()
            got = ar.fused_add_rmsnorm(part.clone(), res, w, 1e-5, odt, twoshot=False)
        for _ in range(3):  # owner-reduces-and-gathers form: same bits as the one-shot form, on every rank
            res2 = res0.clone()
            got2 = ar.fused_add_rmsnorm(part.clone(), res2, w, 1e-5, odt, twoshot=True)
        torch.cuda.synchronize()
        assert got is not None and got2 is not None, (rows, hidden)
        if not (torch.equal(got2, got) and torch.equal(res2, res)):
            ok = False
            print(f"[rank {rank}] two-shot fused add+norm differs from one-shot rows={rows} hidden={hidden}: out max diff "
                  f"{(got2.float() - got.float()).abs().max().item()}, residual equal {torch.equal(res2, res)}", flush=True)
        # push form, row-sharded residual: every normed row = the pull form's bits on every rank; the residual is
        # updated on the rows this rank owns (r % world) and left alone elsewhere
        for _ in range(3):
            res3 = res0.clone()
            got3 = ar.fused_add_rmsnorm_sharded(part.clone(), res3, w, 1e-5, odt)
        torch.cuda.synchronize()
        assert got3 is not None, (rows, hidden)
        mine_rows = torch.arange(rows, device=dev) % world == rank
        if not (torch.equal(got3, got) and torch.equal(res3[mine_rows], res[mine_rows])
                and torch.equal(res3[~mine_rows], res0[~mine_rows])):
            ok = False
            print(f"[rank {rank}] push fused add+norm differs rows={rows} hidden={hidden} f16={f16}: out max diff "
                  f"{(got3.float() - got.float()).abs().max().item()}, owned residual equal "
                  f"{torch.equal(res3[mine_rows], res[mine_rows])}, other rows untouched "
                  f"{torch.equal(res3[~mine_rows], res0[~mine_rows])}", flush=True)
        same_map = hidden in (4096, 8192)
        good = torch.equal(res, res_ref) and (torch.equal(got, want) if same_map else
                                               bool((got.float() - want.float()).abs().max() <= 2.0 ** -6 * want.float().abs().max()))
        if not good:
            ok = False
            print(f"[rank {rank}] fused add+norm differs rows={rows} hidden={hidden} f16={f16}: residual equal "
                  f"{torch.equal(res, res_ref)}, out max diff {(got.float() - want.float()).abs().max().item()}", flush=True)
    parts = [x0.clone() for _ in range(16)]
    resid = x0.clone()
    wn = torch.ones((4096,), device=dev).bfloat16()
    timed(lambda t: ar.fused_add_rmsnorm(t, resid, wn, 1e-5, torch.float16, twoshot=False),
          "nvlink one-shot fused with residual add + RMSNorm")
    timed(lambda t: ar.fused_add_rmsnorm(t, resid, wn, 1e-5, torch.float16, twoshot=True),
          "nvlink two-shot (owner reduces + norms, results gathered) fused with residual add + RMSNorm")

    timed(lambda t: ar.fused_add_rmsnorm_sharded(t, resid, wn, 1e-5, torch.float16),
          "nvlink push (LL packets, row-sharded residual) fused with residual add + RMSNorm")

    def two_step(t):
        ar(t)
        ops.rms_norm(t, wn, 1e-5, residual=resid, out_dtype=torch.float16)

    timed(two_step, "nvlink one-shot, then add + RMSNorm (two launches)")

    # graph-replayed result equals the eager one
    a = x0.clone()
    ar(a)
    b = x0.clone()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        g2 = torch.cuda.CUDAGraph()
        ar(b.clone())
        torch.cuda.synchronize()
        with torch.cuda.graph(g2, stream=s):
            ar(b)
        b.copy_(x0)
        g2.replay()
        torch.cuda.synchronize()
    if not torch.equal(a, b):
        ok = False
        print(f"[rank {rank}] graph replay differs from eager", flush=True)
    # ---- vocab-parallel greedy head: (max, id) exchange vs arg-max of the gathered logits --------------------------
    gen = torch.Generator(device=dev).manual_seed(100)  # same stream on every rank: the full logits are known everywhere

    def first_max(t):  # index of the FIRST maximum per row (mx.argmax / numpy; bf16 rows do repeat their maximum)
        f = t.float()
        cols = torch.arange(f.shape[-1], device=f.device).expand_as(f)
        hit = f == f.max(dim=-1, keepdim=True).values
        return torch.where(hit, cols, torch.full_like(cols, f.shape[-1])).min(dim=-1).values.to(torch.int32)

    shard = 16032  # 128256 / 8
    for it, (rows, dt) in enumerate([(64, torch.bfloat16), (1, torch.bfloat16), (257, torch.float16), (64, torch.bfloat16),
                                     (3, torch.float32), (64, torch.bfloat16), (64, torch.bfloat16)]):
        full = torch.randn((rows, shard * world), device=dev, generator=gen).to(dt)
        if it % 2 == 0:  # exact ties across ranks and inside a slice: the lowest token id must win
            top = full.float().max().item() + 1.0
            for r in range(rows):
                cols = torch.randint(0, shard * world, (3,), device=dev, generator=gen)
                full[r, cols] = top
        want = first_max(full)
        mine = full[:, rank * shard:(rank + 1) * shard]  # strided view: row stride = the full vocab
        got = ar.argmax_vocab_parallel(mine, rank * shard)
        if not torch.equal(got, want):
            ok = False
            bad = (got != want).nonzero().flatten()[:4].tolist()
            print(f"[rank {rank}] vocab-parallel argmax differs (call {it}, rows {rows}, {dt}): rows {bad} got "
                  f"{got[bad].tolist()} want {want[bad].tolist()}", flush=True)
    full = torch.randn((64, shard * world), device=dev, generator=gen).bfloat16()
    mine = full[:, rank * shard:(rank + 1) * shard].contiguous()
    toks = torch.empty((64,), dtype=torch.int32, device=dev)
    timed(lambda t: ar.argmax_vocab_parallel(mine, rank * shard, out=toks),
          "vocab-parallel greedy: slice arg-max + NVLink (max, id) exchange, 64 rows x 16032")
    torch.cuda.synchronize()
    if not torch.equal(toks, first_max(full)):
        ok = False
        print(f"[rank {rank}] vocab-parallel argmax differs after the timed replays", flush=True)

    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("ALLREDUCE " + ("OK" if flag.item() == 1 else "FAILED"), flush=True)
    dist.barrier()
    torch.cuda.synchronize()
    sys.stdout.flush()
    os._exit(0 if flag.item() == 1 else 1)


if __name__ == "__main__":
    main()
