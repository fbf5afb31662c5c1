"""Tensor-parallel sum over NVLink peer memory (kernel: csrc/nvlink_allreduce.cu).

``torch.distributed`` is the plumbing only (rank/world, one ``all_gather_object`` of the CUDA-IPC
handles at construction); the per-layer exchange is our own one-shot kernel through peer-mapped
buffers.  Messages that do not fit the staging buffer or are not 16-byte multiples use
``dist.all_reduce`` (NCCL) -- a size gate, not a fallback for a missing kernel: the constructor
raises if the CUDA library or peer access is unavailable.
"""

from __future__ import annotations

import ctypes
import os

import torch
import torch.distributed as dist

from ..metal import _DTYPE_CODE, get_ops


class NVLinkAllReduce:
    """``ar(x)`` sums ``x`` over the group in place.  Safe inside CUDA-graph capture: pointers are
    fixed at construction and the call epoch lives in device memory."""

    def __init__(self, group=None, device=None, max_bytes: int = 1 << 20, grid: int = 64) -> None:
        if not dist.is_initialized():
            raise RuntimeError("NVLinkAllReduce needs an initialised torch.distributed process group")
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        if self.world not in (2, 4, 8):
            raise ValueError(f"NVLinkAllReduce: world size must be 2, 4 or 8, got {self.world}")
        self.device = torch.device(device if device is not None else torch.cuda.current_device())
        self.ops = get_ops()
        lib = self.lib = self.ops._lib
        lib.b200_ipc_alloc.argtypes = [ctypes.c_int64, ctypes.POINTER(ctypes.c_void_p), ctypes.c_char_p]
        lib.b200_ipc_open.argtypes = [ctypes.c_char_p, ctypes.POINTER(ctypes.c_void_p)]
        lib.b200_ipc_close.argtypes = [ctypes.c_void_p]
        lib.b200_ipc_free.argtypes = [ctypes.c_void_p]
        lib.b200_allreduce_sum.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_int,
                                           ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_void_p),
                                           ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int64,
                                           ctypes.c_int, ctypes.c_void_p]
        lib.b200_allreduce_add_rmsnorm.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                                   ctypes.c_int, ctypes.c_int, ctypes.c_float, ctypes.c_int, ctypes.c_int,
                                                   ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_void_p),
                                                   ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int64,
                                                   ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
        # fused all-reduce+norm: the owner-reduces-and-gathers (two-shot) form from this many ranks up.  Measured SLOWER at
        # every size on one NVSwitch box (2 ranks 18.8 vs 11.4 us, 8 ranks 27.4 vs 23.8: the second flag exchange costs
        # more than the 2.2 MiB of pulls it saves), so it is off by default (profiles/r02_allreduce_tp{2,8}.txt)
        self.twoshot_from = int(os.environ.get("B200_ALLREDUCE_TWOSHOT_FROM", "16"))
        lib.b200_argmax_vocab_parallel.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                                                   ctypes.c_int64, ctypes.c_int, ctypes.c_int,
                                                   ctypes.POINTER(ctypes.c_void_p), ctypes.c_void_p, ctypes.c_int,
                                                   ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
        lib.b200_allreduce_add_rmsnorm_push.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                                        ctypes.c_int, ctypes.c_int, ctypes.c_float, ctypes.c_int,
                                                        ctypes.c_int, ctypes.POINTER(ctypes.c_void_p), ctypes.c_int64,
                                                        ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                                        ctypes.c_void_p]
        # push form with a row-sharded residual stream (fused_add_rmsnorm_sharded); B200_ALLREDUCE_PUSH=0 keeps the
        # decoder on the pull kernel
        self.push = os.environ.get("B200_ALLREDUCE_PUSH", "1") != "0"
        self.half_bytes = (int(max_bytes) + 255) // 256 * 256
        self.grid = int(grid)
        self.max_rows = 1024  # rows of one vocab-parallel arg-max call (slot table: 2 parities x world x rows x 16 B)
        with torch.cuda.device(self.device):
            # [2 halves of partials][2 x 2 halves of two-shot results]; [grid][world] flags x 2 exchanges
            self._own_symm, h_symm = self._alloc(6 * self.half_bytes)
            self._own_flags, h_flags = self._alloc(2 * self.grid * self.world * 4)
            self._own_slots, h_slots = self._alloc(2 * self.world * self.max_rows * 16)
            # packets are 2x the payload; [2 parities][inbox | outbox], rows rounded up to the world size
            self.box_bytes = 4 * (2 * self.half_bytes + self.world * 8192 * 4 + 256)
            self._own_box, h_box = self._alloc(self.box_bytes)
            self.push_epoch = torch.zeros((self.grid,), dtype=torch.int32, device=self.device)
            self.epoch = torch.zeros((self.grid,), dtype=torch.int32, device=self.device)
            self.argmax_epoch = torch.zeros((self.max_rows,), dtype=torch.int32, device=self.device)
            handles: list = [None] * self.world
            dist.all_gather_object(handles, (h_symm, h_flags, h_slots, h_box), group=group)
            self._opened: list[int] = []
            symm, flags, slots, boxes = [], [], [], []
            for r, (hs, hf, hl, hb) in enumerate(handles):
                if r == self.rank:
                    symm.append(self._own_symm)
                    flags.append(self._own_flags)
                    slots.append(self._own_slots)
                    boxes.append(self._own_box)
                else:
                    symm.append(self._open(hs))
                    flags.append(self._open(hf))
                    slots.append(self._open(hl))
                    boxes.append(self._open(hb))
            self._boxes = (ctypes.c_void_p * self.world)(*boxes)
            self._symm = (ctypes.c_void_p * self.world)(*symm)
            self._flags = (ctypes.c_void_p * self.world)(*flags)
            self._slots = (ctypes.c_void_p * self.world)(*slots)
            torch.cuda.synchronize(self.device)
        dist.barrier(group=group)  # every rank has zeroed and mapped everything before the first call
        self.calls = 0
        self.nccl_calls = 0

    def _alloc(self, nbytes: int) -> tuple[int, bytes]:
        ptr = ctypes.c_void_p()
        handle = ctypes.create_string_buffer(64)
        self.ops._check(self.lib.b200_ipc_alloc(nbytes, ctypes.byref(ptr), handle))
        return int(ptr.value), handle.raw

    def _open(self, handle: bytes) -> int:
        ptr = ctypes.c_void_p()
        self.ops._check(self.lib.b200_ipc_open(handle, ctypes.byref(ptr)))
        self._opened.append(int(ptr.value))
        return int(ptr.value)

    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        nbytes = x.numel() * x.element_size()
        if (not x.is_contiguous()) or nbytes % 16 or x.data_ptr() % 16 or nbytes > self.half_bytes \
                or x.dtype not in _DTYPE_CODE:
            dist.all_reduce(x, group=self.group)
            self.nccl_calls += 1
            return x
        rc = self.lib.b200_allreduce_sum(x.data_ptr(), x.numel(), _DTYPE_CODE[x.dtype], self._symm, self._flags,
                                         self.epoch.data_ptr(), self.rank, self.world, self.half_bytes, self.grid,
                                         torch.cuda.current_stream().cuda_stream)
        self.ops._check(rc)
        self.ops.launches += 1
        self.calls += 1
        return x

    def fused_add_rmsnorm(self, x: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor | None, eps: float,
                          out_dtype: torch.dtype | None = None, twoshot: bool | None = None) -> torch.Tensor | None:
        """``residual += allreduce(x); return rms_norm(residual) * weight`` in one kernel (x is left holding this
        rank's partial).  Returns ``None`` when the shape does not qualify -- the caller then runs the two-step
        sequence, which gives the same bits."""
        hidden = x.shape[-1]
        x2, r2 = x.reshape(-1, hidden), residual.reshape(-1, hidden)
        nbytes = x2.numel() * 2
        if x2.shape[0] > 4 * self.grid:  # prefill-sized: the norm wants one CTA per row, not grid CTAs looping over rows
            return None
        if (x.dtype not in (torch.bfloat16, torch.float16) or not x2.is_contiguous() or not r2.is_contiguous()
                or hidden % 8 or hidden > 8192 or nbytes > self.half_bytes or x2.data_ptr() % 16 or r2.data_ptr() % 16
                or (weight is not None and (weight.dtype != x.dtype or weight.data_ptr() % 16))):
            return None
        out = torch.empty(x2.shape, dtype=out_dtype or x.dtype, device=x.device)
        rc = self.lib.b200_allreduce_add_rmsnorm(
            x2.data_ptr(), r2.data_ptr(), None if weight is None else weight.data_ptr(), out.data_ptr(), x2.shape[0],
            hidden, float(eps), _DTYPE_CODE[x.dtype], _DTYPE_CODE[out.dtype], self._symm, self._flags,
            self.epoch.data_ptr(), self.rank, self.world, self.half_bytes, self.grid,
            int(self.world >= self.twoshot_from if twoshot is None else twoshot), torch.cuda.current_stream().cuda_stream)
        self.ops._check(rc)
        self.ops.launches += 1
        self.calls += 1
        return out.reshape(x.shape)

    def fused_add_rmsnorm_sharded(self, x: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor | None, eps: float,
                                  out_dtype: torch.dtype | None = None) -> torch.Tensor | None:
        """Push form of ``fused_add_rmsnorm`` with a ROW-SHARDED residual stream: returns ``rms_norm(residual +
        allreduce(x)) * weight`` for every row, the same bits as ``fused_add_rmsnorm``, but ``residual`` is read and
        updated only on the rows this rank owns (row r -> rank r % world); the other rows are left alone.  A forward
        pass must use this form for all of its sums or for none.  ``None`` when the shape does not qualify."""
        hidden = x.shape[-1]
        x2, r2 = x.reshape(-1, hidden), residual.reshape(-1, hidden)
        rows = x2.shape[0]
        if not self.push or self.grid % self.world or rows > 4 * self.grid:
            return None
        packet_rows = (rows + self.world - 1) // self.world * self.world
        if (x.dtype not in (torch.bfloat16, torch.float16) or not x2.is_contiguous() or not r2.is_contiguous()
                or hidden % 8 or hidden > 8192 or packet_rows * hidden * 4 > (self.box_bytes // 4) // 256 * 256
                or x2.data_ptr() % 16 or r2.data_ptr() % 16
                or (weight is not None and (weight.dtype != x.dtype or weight.data_ptr() % 16))):
            return None
        out = torch.empty(x2.shape, dtype=out_dtype or x.dtype, device=x.device)
        rc = self.lib.b200_allreduce_add_rmsnorm_push(
            x2.data_ptr(), r2.data_ptr(), None if weight is None else weight.data_ptr(), out.data_ptr(), rows, hidden,
            float(eps), _DTYPE_CODE[x.dtype], _DTYPE_CODE[out.dtype], self._boxes, self.box_bytes,
            self.push_epoch.data_ptr(), self.rank, self.world, self.grid, torch.cuda.current_stream().cuda_stream)
        self.ops._check(rc)
        self.ops.launches += 1
        self.calls += 1
        return out.reshape(x.shape)

    def argmax_vocab_parallel(self, logits: torch.Tensor, vocab_offset: int, out: torch.Tensor | None = None) -> torch.Tensor:
        """Greedy tokens of a vocab-parallel head: ``logits`` [rows, vocab / world] is this rank's slice starting at
        token ``vocab_offset``; returns int32 [rows], the arg-max over every rank's slice (lowest id on ties), the
        same on all ranks.  Two launches: arg-max of the slice, then a (value, id) exchange pushed into the peers'
        slot tables over NVLink."""
        x2 = logits.reshape(-1, logits.shape[-1])
        if x2.stride(1) != 1:
            x2 = x2.contiguous()
        rows = x2.shape[0]
        if rows > self.max_rows:
            raise ValueError(f"argmax_vocab_parallel: {rows} rows exceed the slot table ({self.max_rows})")
        if out is None:
            out = torch.empty((rows,), dtype=torch.int32, device=x2.device)
        elif out.dtype != torch.int32 or out.numel() != rows or not out.is_contiguous():
            raise ValueError("argmax_vocab_parallel: out must be a contiguous int32 [rows]")
        rc = self.lib.b200_argmax_vocab_parallel(
            x2.data_ptr(), out.data_ptr(), rows, x2.shape[1], x2.stride(0), _DTYPE_CODE[x2.dtype], int(vocab_offset),
            self._slots, self.argmax_epoch.data_ptr(), self.rank, self.world, self.max_rows,
            torch.cuda.current_stream().cuda_stream)
        self.ops._check(rc)
        self.ops.launches += 2
        return out

    def close(self) -> None:
        torch.cuda.synchronize(self.device)
        for p in self._opened:
            self.lib.b200_ipc_close(p)
        self._opened = []
        for p in (self._own_symm, self._own_flags, self._own_slots, self._own_box):
            if p:
                self.lib.b200_ipc_free(p)
        self._own_symm = self._own_flags = self._own_slots = self._own_box = 0
