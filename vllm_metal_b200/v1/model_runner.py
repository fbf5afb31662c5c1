"""Paged-generate runner: packs a step, runs the model, returns greedy tokens.

Covers the paged path of the reference's ``MetalModelRunner``
(vllm_metal/v1/model_runner.py: ``_start_paged_forward`` :1138-1410 packs decode
rows first then prefill chunks and calls ``prepare_grouped``; ``sample_tokens``
:2733-2783 takes the last row of every segment; greedy = ``mlx_greedy_tokens``,
v1/sampling_batch.py:342-348).  B200 design: all per-step integer metadata
travels in ONE pinned staging buffer -> one H2D copy into persistent device
buffers, the decode forward is a CUDA graph replayed per step (captured per
batch size and split-KV bucket), and the only D2H is the sampled token ids.
The reference's deferred one-step-ahead sampling (v1/decode_pipeline.py) is
replaced by stream order + the graph.
"""

from __future__ import annotations

from collections.abc import Sequence

import numpy as np
import torch

from ..attention.context import clear_context, get_context, prepare_grouped, prepare_unified
from ..attention.impls.sdpa import _KernelMetadata, _kernel_metadata
from ..metal import get_ops
from ..metal.constants import PARTITION_SIZE


class StepBuffers:
    """Persistent device metadata for a fixed-capacity decode batch + its pinned mirror."""

    def __init__(self, max_batch: int, max_blocks: int, device) -> None:
        self.max_batch, self.max_blocks = max_batch, max_blocks
        b = max_batch
        # layout in int32 words: slot_mapping (int64 -> 2b words, first for alignment)
        sizes = {"slot_mapping": 2 * b, "input_ids": b, "positions": b, "seq_lens": b,
                 "cu_seqlens": b + 2, "block_tables": b * max_blocks}
        self.offsets, off = {}, 0
        for k, n in sizes.items():
            self.offsets[k] = (off, n)
            off += n + (n & 1)  # keep every field 8-byte aligned
        self.words = off
        self.host = torch.zeros(off, dtype=torch.int32).pin_memory()
        self.dev = torch.zeros(off, dtype=torch.int32, device=device)
        self.host_np = self.host.numpy()
        self.out_tokens = torch.zeros(b, dtype=torch.int32, device=device)
        self.out_host = torch.zeros(b, dtype=torch.int32).pin_memory()
        # one-step-ahead pipeline (submit_decode): the previous step's sampled tokens stay on the device as the next
        # step's inputs; two pinned read-back buffers alternate so a step's tokens can be read after the next submit
        self.prev_tokens = torch.zeros(b, dtype=torch.int32, device=device)
        self.out_host2 = [torch.zeros(b, dtype=torch.int32).pin_memory() for _ in range(2)]
        self.gather_host = torch.zeros(2 * b, dtype=torch.int64).pin_memory()
        self.gather_dev = torch.zeros(2 * b, dtype=torch.int64, device=device)

    def view(self, name: str, on_host: bool = False):
        off, n = self.offsets[name]
        src = self.host if on_host else self.dev
        v = src[off:off + n]
        if name == "slot_mapping":
            return v.view(torch.int64)
        if name == "cu_seqlens":
            return v[: self.max_batch + 1]
        if name == "block_tables":
            return v.view(self.max_batch, self.max_blocks)
        return v

    def np_view(self, name: str):
        off, n = self.offsets[name]
        v = self.host_np[off:off + n]
        if name == "slot_mapping":
            return v.view(np.int64)
        if name == "cu_seqlens":
            return v[: self.max_batch + 1]
        if name == "block_tables":
            return v.reshape(self.max_batch, self.max_blocks)
        return v

    def upload(self) -> int:
        self.dev.copy_(self.host, non_blocking=True)
        return self.words * 4


class _PendingTokens:
    """Tokens of a submitted decode step; ``tokens()`` waits for the read-back."""

    def __init__(self, event, host: torch.Tensor, n: int) -> None:
        self._event, self._host, self._n = event, host, n
        self._tokens: list[int] | None = None

    def tokens(self) -> list[int]:
        if self._tokens is None:
            self._event.synchronize()
            self._tokens = self._host[: self._n].tolist()
        return self._tokens


class PagedModelRunner:
    """Greedy generation over a paged KV cache (single KV group)."""

    def __init__(self, model, kv_cache, *, max_batch: int = 64, max_model_len: int = 8192,
                 use_cuda_graph: bool = True) -> None:
        self.model, self.kv_cache = model, kv_cache
        self.device = kv_cache.device
        self.block_size = kv_cache.block_size
        self.max_batch = max_batch
        self.max_blocks = (max_model_len + self.block_size - 1) // self.block_size
        # the MoE expert path reads its per-expert row counts back to the host every block: not capturable
        has_moe = any(hasattr(getattr(layer, "mlp", None), "block") for layer in getattr(model, "layers", []))
        self.use_cuda_graph = use_cuda_graph and not has_moe
        self.group_block_sizes: tuple[int, ...] | None = None  # set by the worker for multi-group KV layouts
        self.buf = StepBuffers(max_batch, self.max_blocks, self.device)
        self._graphs: dict[tuple[int, int, bool], torch.cuda.CUDAGraph] = {}
        # kernels of ours inside each captured graph (bench.py's gpu_launches)
        self.launches_per_graph: dict[tuple[int, int, bool], int] = {}
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    def _greedy(self, logits, out=None):
        greedy = getattr(self.model, "greedy", None)  # vocab-parallel heads exchange (max, id) between the ranks
        return greedy(logits, out=out) if greedy is not None else get_ops().argmax(logits, out=out)

    # -- generic (eager) path: any mix of prefill chunks and decode rows -----------
    def execute(self, input_ids: Sequence[int], decode_requests, prefill_requests) -> list[int]:
        """One packed forward; returns the greedy token of every segment's last row.

        ``decode_requests``: ``(block_ids, seq_len)``; ``prefill_requests``:
        ``(block_ids, num_tokens, start_pos)`` -- the ``prepare_unified`` contract.
        """
        if self.group_block_sizes is not None:
            # mixed full / sliding-window KV groups: every request carries one block-id list per scheduler group
            prepare_grouped(decode_requests, prefill_requests, self.group_block_sizes)
        else:
            prepare_unified(decode_requests, prefill_requests, self.block_size)
        ctx = get_context()
        try:
            ids = torch.tensor(list(input_ids), dtype=torch.int32).pin_memory().to(self.device, non_blocking=True)
            last_rows = torch.tensor([c - 1 for c in ctx.cu_seqlens[1:]], dtype=torch.int64).pin_memory().to(
                self.device, non_blocking=True)
            logits = self.model(ids, ctx, self.kv_cache, logits_rows=last_rows)
            toks = self._greedy(logits)
            out = toks.cpu().tolist()
            # bytes that crossed PCIe for this step: ids + logit rows + the kernel metadata sdpa uploads once
            # (slot_mapping int64 [T], seq_lens / cu_seqlens int32, block tables int32), token ids back
            nseq = len(ctx.cu_seqlens) - 1
            maxb = max((len(t) for t in (ctx.block_tables or [])), default=0)
            self.h2d_bytes += ids.numel() * 4 + last_rows.numel() * 8 + len(ctx.slot_mapping) * 8 + \
                (2 * nseq + 1) * 4 + nseq * maxb * 4 + len(ctx.slot_mapping) * 4
            self.d2h_bytes += len(out) * 4
            return out
        finally:
            clear_context()

    # -- speculative verification: K+1-row decode windows ------------------------------
    # Window-mode layout (one cu_seqlens segment per request; reference runner property ``merge_verify_windows``,
    # attention/context.py:187-197): the decode kernel then shares every KV read between a window's rows.  False
    # gives the expanded one-row-per-segment layout -- the same numbers bit for bit, K+1 times the KV traffic.
    merge_verify_windows: bool = True

    def verify(self, input_ids: Sequence[int], decode_requests) -> list[list[int]]:
        """One verification step.  ``decode_requests``: ``(block_ids, seq_len, num_tokens)`` with ``seq_len`` tokens
        already cached and ``num_tokens`` = 1 + drafts; ``input_ids``: the packed window tokens (last accepted token,
        then the drafts).  Returns the greedy token after EVERY window row, one list per request: row j's token
        verifies draft j+1 and the first mismatch (or the last row) yields the bonus token."""
        reqs = [(r[0], r[1], r[2] if len(r) > 2 else 1) for r in decode_requests]
        if self.group_block_sizes is not None:
            prepare_grouped(reqs, [], self.group_block_sizes, merge_verify_windows=self.merge_verify_windows)
        else:
            prepare_unified(reqs, [], self.block_size, merge_verify_windows=self.merge_verify_windows)
        ctx = get_context()
        try:
            ids = torch.tensor(list(input_ids), dtype=torch.int32).pin_memory().to(self.device, non_blocking=True)
            self.last_verify_logits = self.model(ids, ctx, self.kv_cache, logits_rows=None)
            toks = self._greedy(self.last_verify_logits).cpu().tolist()
            nseq = len(ctx.cu_seqlens) - 1
            maxb = max((len(t) for t in (ctx.block_tables or [])), default=0)
            self.h2d_bytes += ids.numel() * 4 + len(ctx.slot_mapping) * 12 + (2 * nseq + 1) * 4 + nseq * maxb * 4
            self.d2h_bytes += len(toks) * 4
            out, at = [], 0
            for r in reqs:
                out.append(toks[at:at + r[2]])
                at += r[2]
            return out
        finally:
            clear_context()

    # -- decode fast path: fixed batch, persistent buffers, CUDA graph ---------------
    def _meta(self, batch: int, max_seq_len: int) -> _KernelMetadata:
        b = self.buf
        return _KernelMetadata(
            slot_mapping=b.view("slot_mapping")[:batch],
            seq_lens=b.view("seq_lens")[:batch],
            cu_seqlens_q=b.view("cu_seqlens")[: batch + 1],
            block_tables=b.view("block_tables")[:batch],
            block_size=self.block_size,
            max_seq_len=max_seq_len,
            positions=b.view("positions")[:batch],
        )

    def _forward_decode(self, batch: int, max_seq_len: int, advance: bool = False) -> None:
        meta = self._meta(batch, max_seq_len)
        logits = self.model(self.buf.view("input_ids")[:batch], None, self.kv_cache, meta=meta)
        self._greedy(logits, out=self.buf.out_tokens[:batch])
        if advance:
            self.advance_on_device(batch)

    def _bucket(self, max_seq_len: int) -> int:
        """Graphs are keyed by the split-KV bucket: max_seq_len rounded up to a
        multiple of PARTITION_SIZE (the partition plan depends only on it)."""
        return (max_seq_len + PARTITION_SIZE - 1) // PARTITION_SIZE * PARTITION_SIZE

    def fill_decode(self, token_ids: Sequence[int], block_ids: Sequence[Sequence[int]],
                    seq_lens: Sequence[int]) -> int:
        """Write one decode step's metadata into the pinned mirror (no GPU work).
        ``seq_lens`` = tokens already cached per request.  Returns max KV length."""
        b, n = self.buf, len(token_ids)
        pos = np.asarray(seq_lens, dtype=np.int64)
        tables = b.np_view("block_tables")
        if isinstance(block_ids, np.ndarray):
            tables[:n, : block_ids.shape[1]] = block_ids
        else:
            for i, ids in enumerate(block_ids):
                tables[i, : len(ids)] = ids
        blk = tables[np.arange(n), pos // self.block_size].astype(np.int64)
        b.np_view("slot_mapping")[:n] = blk * self.block_size + pos % self.block_size
        b.np_view("input_ids")[:n] = np.asarray(token_ids, dtype=np.int32)
        b.np_view("positions")[:n] = pos
        b.np_view("seq_lens")[:n] = pos + 1
        b.np_view("cu_seqlens")[: n + 1] = np.arange(n + 1, dtype=np.int32)
        return int(pos.max()) + 1

    def decode_step(self, token_ids, block_ids, seq_lens, *, sync: bool = True):
        """One pure-decode step through the public runner API: host metadata ->
        pinned -> H2D -> (graph) forward -> D2H token ids."""
        n = len(token_ids)
        max_len = self.fill_decode(token_ids, block_ids, seq_lens)
        self.h2d_bytes += self.buf.upload()
        self.run_decode(n, max_len)
        self.buf.out_host[:n].copy_(self.buf.out_tokens[:n], non_blocking=True)
        self.d2h_bytes += n * 4
        if sync:
            torch.cuda.current_stream().synchronize()
            return self.buf.out_host[:n].tolist()
        return None

    def submit_decode(self, token_ids, block_ids, seq_lens, pending_rows: Sequence[int] | None = None):
        """One pure-decode step WITHOUT waiting for its tokens (one-step-ahead pipelining, reference
        v1/decode_pipeline.py:1-26): host metadata -> pinned -> H2D -> graph -> async D2H, returns a handle whose
        ``tokens()`` blocks on the read-back event.  ``pending_rows[i] >= 0`` says request i's input token is row
        ``pending_rows[i]`` of the PREVIOUS submitted step, whose value the host has not seen: it is gathered on the
        device (``token_ids[i]`` is ignored); -1 = ``token_ids[i]`` is the real token."""
        n = len(token_ids)
        b = self.buf
        gather = [(i, r) for i, r in enumerate(pending_rows or ()) if r >= 0]
        ids = [0 if (pending_rows and pending_rows[i] >= 0) else int(t) for i, t in enumerate(token_ids)]
        # the pinned mirrors are reused every step: the previous step's H2D copies (queued behind the step before it)
        # must have left them before the host writes this step's metadata
        prev_upload = getattr(self, "_upload_event", None)
        if prev_upload is not None:
            prev_upload.synchronize()
        max_len = self.fill_decode(ids, block_ids, seq_lens)
        self.h2d_bytes += b.upload()
        if gather:
            # prev_tokens holds the previous step's argmax (copied before this step's argmax overwrites out_tokens)
            g = len(gather)
            b.gather_host[:g] = torch.tensor([i for i, _ in gather], dtype=torch.int64)
            b.gather_host[g:2 * g] = torch.tensor([r for _, r in gather], dtype=torch.int64)
            b.gather_dev[:2 * g].copy_(b.gather_host[:2 * g], non_blocking=True)
            self.h2d_bytes += 16 * g
            b.view("input_ids").index_copy_(0, b.gather_dev[:g], b.prev_tokens.index_select(0, b.gather_dev[g:2 * g]))
        self._upload_event = torch.cuda.Event()
        self._upload_event.record()
        self.run_decode(n, max_len)
        b.prev_tokens[:n].copy_(b.out_tokens[:n])
        self._submit_serial = getattr(self, "_submit_serial", 0) + 1
        host = b.out_host2[self._submit_serial & 1]
        host[:n].copy_(b.out_tokens[:n], non_blocking=True)
        self.d2h_bytes += n * 4
        ev = torch.cuda.Event()
        ev.record()
        return _PendingTokens(ev, host, n)

    def run_decode(self, batch: int, max_seq_len: int, *, advance: bool = False) -> None:
        """Forward + argmax over the persistent buffers (graph replay when enabled).
        ``advance=True`` appends the device-side bookkeeping kernel so steps can be
        replayed back to back with no host in the loop."""
        if not self.use_cuda_graph:
            self._forward_decode(batch, max_seq_len, advance)
            return
        key = (batch, self._bucket(max_seq_len), advance)
        g = self._graphs.get(key)
        if g is None:
            if advance:
                snap = self.buf.dev.clone()
            # warm up on a side stream (lazy module load, workspace growth), then capture
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                self._forward_decode(batch, key[1], advance)
            torch.cuda.current_stream().wait_stream(s)
            torch.cuda.synchronize()
            if advance:
                self.buf.dev.copy_(snap)  # the warm-up run moved the requests one slot on
            ops = get_ops()
            n0 = ops.launches
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._forward_decode(batch, key[1], advance)
            self.launches_per_graph[key] = ops.launches - n0
            self._graphs[key] = g
        g.replay()

    def advance_on_device(self, batch: int) -> None:
        """Device-side bookkeeping for back-to-back decode steps with no host in
        the loop: feed the sampled tokens back and move every request one slot on
        (one kernel; the reference does this on the host, model_runner.py:1200-1240)."""
        b = self.buf
        get_ops().advance_decode(b.view("input_ids")[:batch], b.view("positions")[:batch],
                                 b.view("seq_lens")[:batch], b.view("slot_mapping")[:batch],
                                 b.view("block_tables")[:batch], b.out_tokens[:batch], self.block_size)


__all__ = ["PagedModelRunner", "StepBuffers", "_kernel_metadata"]
