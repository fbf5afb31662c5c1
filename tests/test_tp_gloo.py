"""Tensor-parallel host logic on CPU with a real world-size-2 ``gloo`` group.

Two processes shard one quantized attention/MLP block the way ``lm/decoder.py`` does
(column-parallel q/k/v + gate/up via ``shard_rows``, row-parallel o_proj/down via ``shard_cols``),
compute their partial products with the *oracle's* dequantisation (no CUDA here), sum them with
``dist.all_reduce`` and compare with the unsharded computation.  This pins the sharding
arithmetic -- packed words, scales and biases cut on the same group boundaries, dense bias on rank
0 only, attention heads split without replication -- that the GPU path reuses unchanged.
"""

from __future__ import annotations

import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import affine_quant as oq


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _dense(m) -> np.ndarray:
    """fp32 dequantised weight of a QuantizedLinear via the oracle (bit-exact spec)."""
    return oq.dequantize(m.weight.numpy().view(np.uint32), m.scales.float().numpy(), m.biases.float().numpy(),
                         m.group_size, m.bits).astype(np.float64)


def _worker(rank: int, world: int, port: int, bits: int, group: int, out: dict) -> None:
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from vllm_metal_b200.quant.linear import QuantizedLinear

        g = torch.Generator().manual_seed(7)  # same weights on every rank, like a checkpoint
        hidden, inter, m_rows = 256, 512, 5
        w_up = torch.randn((inter, hidden), generator=g) * 0.05
        w_down = torch.randn((hidden, inter), generator=g) * 0.05
        b_down = torch.randn((hidden,), generator=g) * 0.1
        x = torch.randn((m_rows, hidden), generator=g).double().numpy()
        up = QuantizedLinear.from_dense(w_up, None, group, bits, torch.bfloat16)
        down = QuantizedLinear.from_dense(w_down, b_down, group, bits, torch.bfloat16)
        full = (x @ _dense(up).T) @ _dense(down).T + down.bias.double().numpy()

        up_r, down_r = up.shard_rows(rank, world), down.shard_cols(rank, world)
        assert up_r.out_features == inter // world and down_r.in_features == inter // world
        assert (down_r.bias is not None) == (rank == 0)  # dense bias added once
        # the shards are literal slices of the packed tensors: no requantisation
        lo, hi = rank * inter // world, (rank + 1) * inter // world
        assert torch.equal(up_r.weight, up.weight[lo:hi]) and torch.equal(up_r.scales, up.scales[lo:hi])
        assert torch.equal(down_r.scales, down.scales[:, lo // group:hi // group])
        h = x @ _dense(up_r).T                       # [M, inter/world]: this rank's columns
        part = h @ _dense(down_r).T                  # partial sum over this rank's K slice
        if down_r.bias is not None:
            part = part + down_r.bias.double().numpy()
        t = torch.from_numpy(part)
        dist.all_reduce(t)                           # the exchange step of the path
        np.testing.assert_allclose(t.numpy(), full, rtol=1e-10, atol=1e-10)
        gathered = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(gathered, t)
        assert all(torch.equal(gathered[0], o) for o in gathered)

        # heads: kv heads split without replication, q heads follow their kv group
        n_heads, n_kv, d = 8, 4, 32
        wq = QuantizedLinear.from_dense(torch.randn((n_heads * d, hidden), generator=g) * 0.05, None, group, bits)
        wq_r = wq.shard_rows(rank, world)
        per = n_heads // world
        np.testing.assert_array_equal(_dense(wq_r), _dense(wq)[rank * per * d:(rank + 1) * per * d])
        assert (n_kv // world) * (n_heads // n_kv) == per
        with pytest.raises(ValueError):
            QuantizedLinear.from_dense(torch.randn((64, 192), generator=g), None, 64, bits).shard_cols(rank, world)

        # vocab-parallel head: each rank's logits slice -> (max, global id) -> exchange -> first maximum overall
        # (csrc/nvlink_allreduce.cu argmax_exchange_kernel does this step over NVLink; the arithmetic is pinned here)
        vocab = 96
        head = QuantizedLinear.from_dense(torch.randn((vocab, hidden), generator=g) * 0.05, None, group, bits)
        head_r = head.shard_rows(rank, world)
        logits = torch.from_numpy(x @ _dense(head).T).bfloat16().double()  # bf16 rows repeat their maximum
        logits[1, [5, 70]] = logits.max() + 1.0                             # an exact tie across the two ranks
        logits[2, [60, 61]] = logits.max() + 1.0                            # and one inside rank 1's slice
        n = vocab // world
        mine = logits[:, rank * n:(rank + 1) * n]
        np.testing.assert_array_equal(_dense(head_r), _dense(head)[rank * n:(rank + 1) * n])
        val, idx = mine.max(dim=-1).values, torch.from_numpy(np.argmax(mine.numpy(), axis=-1)) + rank * n
        pair = torch.stack([val, idx.double()], dim=-1)
        pairs = [torch.empty_like(pair) for _ in range(world)]
        dist.all_gather(pairs, pair)
        best_v, best_i = pairs[0][:, 0].clone(), pairs[0][:, 1].clone()
        for pr in pairs[1:]:
            take = (pr[:, 0] > best_v) | ((pr[:, 0] == best_v) & (pr[:, 1] < best_i))
            best_v, best_i = torch.where(take, pr[:, 0], best_v), torch.where(take, pr[:, 1], best_i)
        np.testing.assert_array_equal(best_i.long().numpy(), np.argmax(logits.numpy(), axis=-1))  # numpy: first maximum
        assert int(best_i[1]) == 5 and int(best_i[2]) == 60
        out[rank] = "ok"
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("bits,group", [(4, 64), (8, 32)])
def test_world2_sharded_block_equals_unsharded(bits, group):
    world, port = 2, _free_port()
    with mp.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_worker, args=(world, port, bits, group, out), nprocs=world, join=True)
        assert dict(out) == {0: "ok", 1: "ok"}
