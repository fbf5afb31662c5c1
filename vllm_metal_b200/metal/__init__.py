"""Native operator bridge: ``get_ops()`` loads the sm_100a C-ABI library.

Reference counterpart: ``vllm_metal/metal/__init__.py:195-301`` (``get_ops``
imports the nanobind module ``_paged_ops`` and initialises the Metal shader
libraries).  Here the same entry point ``dlopen``s ``lib_b200_paged_ops.so``
(``include/b200_paged_ops.h``) through ctypes and returns an object with the
reference's operator names, argument order and error behaviour
(``paged_ops.cpp:1718-1749,1790-1854``) taking ``torch`` CUDA tensors:

* ``paged_attention_primitive(query, key_cache, value_cache, num_kv_heads,
  scale, softcap, block_tables, seq_lens, cu_seqlens_q, block_size,
  max_seq_len, sliding_window, out, ..., window_seqlen_q=1, sinks=None)``
* ``reshape_and_cache(key, value, key_cache, value_cache, slot_mapping)``
* ``min_decode_grid()``, ``PARTITION_SIZE``

plus the quantized-linear calls the reference makes into MLX
(``quantized_matmul``, ``dequantize``).  There is no CPU or eager fallback: a
missing or stale library, a CPU tensor or an unsupported dtype raises.
"""

from __future__ import annotations

import ctypes
import logging
import os
from typing import Any

import torch

from . import constants
from .constants import DTYPE_BF16, DTYPE_F16, DTYPE_F32, PA_WINDOW_MAX_HEAD_SIZE as PA_WINDOW_MAX_HEAD, PARTITION_SIZE

logger = logging.getLogger(__name__)

_ops: _Ops | None = None

_DTYPE_CODE = {torch.float32: DTYPE_F32, torch.float16: DTYPE_F16, torch.bfloat16: DTYPE_BF16}

_ERR_INVALID_ARGUMENT = -1
_ERR_RUNTIME = -2
_ERR_UNSUPPORTED = -3


class PagedAttnArgs(ctypes.Structure):
    """Mirror of ``b200_paged_attn_args`` (include/b200_paged_ops.h)."""

    _fields_ = [
        ("out", ctypes.c_void_p),
        ("query", ctypes.c_void_p),
        ("key_cache", ctypes.c_void_p),
        ("value_cache", ctypes.c_void_p),
        ("block_tables", ctypes.c_void_p),
        ("seq_lens", ctypes.c_void_p),
        ("cu_seqlens_q", ctypes.c_void_p),
        ("sinks", ctypes.c_void_p),
        ("workspace", ctypes.c_void_p),
        ("workspace_bytes", ctypes.c_size_t),
        ("num_q_tokens", ctypes.c_int32),
        ("num_seqs", ctypes.c_int32),
        ("num_q_heads", ctypes.c_int32),
        ("num_kv_heads", ctypes.c_int32),
        ("head_dim", ctypes.c_int32),
        ("block_size", ctypes.c_int32),
        ("max_blocks_per_seq", ctypes.c_int32),
        ("max_seq_len", ctypes.c_int32),
        ("sliding_window", ctypes.c_int32),
        ("dtype", ctypes.c_int32),
        ("force_partitions", ctypes.c_int32),
        ("kernel", ctypes.c_int32),
        ("q_stride", ctypes.c_int64),
        ("out_stride", ctypes.c_int64),
        ("kv_block_stride", ctypes.c_int64),
        ("kv_token_stride", ctypes.c_int64),
        ("kv_head_stride", ctypes.c_int64),
        ("num_blocks", ctypes.c_int64),
        ("scale", ctypes.c_float),
        ("softcap", ctypes.c_float),
        ("out_dtype", ctypes.c_int32),
        ("window_seqlen_q", ctypes.c_int32),
    ]


# Every symbol include/*.h declares; the CPU test-suite checks the library
# exports all of them (tests/test_host_logic.py::test_library_exports_every_declared_symbol).
EXPORTED_SYMBOLS = (
    "b200_last_error",
    "b200_abi_version",
    "b200_set_pdl",
    "b200_partition_size",
    "b200_min_decode_grid",
    "b200_sm_count",
    "b200_reshape_and_cache",
    "b200_copy_blocks",
    "b200_paged_attention_workspace_bytes",
    "b200_paged_attention_num_launches",
    "b200_paged_attention",
    "b200_paged_attention_static_scratch",
    "b200_qlinear",
    "b200_qlinear_experts",
    "b200_qlinear_prepack",
    "b200_qlinear_prepacked",
    "b200_set_qlinear_variant",
    "b200_set_qlinear_trace",
    "b200_set_qlinear_kernel",
    "b200_set_qlinear_tuning",
    "b200_dequantize_rows",
    "b200_qstream_blob_bytes",
    "b200_qstream_workspace_bytes",
    "b200_qstream_prepack",
    "b200_qstream_unpack",
    "b200_qstream_linear",
    "b200_cast_bf16_to_f16",
    "b200_set_qstream_grid",
    "b200_set_qstream_trace",
    "b200_rms_norm",
    "b200_rope",
    "b200_norm_rope_cache",
    "b200_silu_mul",
    "b200_argmax",
    "b200_add_inplace",
    "b200_advance_decode",
    "b200_ipc_alloc",
    "b200_ipc_open",
    "b200_ipc_close",
    "b200_ipc_free",
    "b200_allreduce_sum",
    "b200_allreduce_add_rmsnorm",
    "b200_allreduce_add_rmsnorm_push",
    "b200_argmax_vocab_parallel",
)


def library_path() -> str:
    from .build import output_path

    return str(output_path())


def load_library() -> ctypes.CDLL:
    """dlopen the prebuilt library; raise (never fall back) when it is missing
    or stale (reference: metal/__init__.py:225-250)."""
    from .build import build, output_path, stale_artifacts

    if os.environ.get("VLLM_METAL_BUILD_FROM_SOURCE", "0") not in ("", "0"):
        build()
    so = output_path()
    override = os.environ.get("B200_PAGED_OPS_LIB")  # tuning experiments only (profiles/build_variants.sh)
    if override:
        from pathlib import Path

        so = Path(override)
    if not so.exists():
        raise RuntimeError(
            f"Prebuilt native extension not found at {so}. Run "
            f"`python -m vllm_metal_b200.metal.build` (needs nvcc) or set "
            f"VLLM_METAL_BUILD_FROM_SOURCE=1; there is no CPU fallback."
        )
    stale = [] if override else stale_artifacts()
    if stale:
        raise RuntimeError(
            f"Prebuilt CUDA artifacts are stale vs the current sources "
            f"({', '.join(p.name for p in stale)}): a kernel source was edited without "
            f"rebuilding. Run `python -m vllm_metal_b200.metal.build`."
        )
    lib = ctypes.CDLL(str(so))
    lib.b200_last_error.restype = ctypes.c_char_p
    lib.b200_paged_attention_workspace_bytes.restype = ctypes.c_size_t
    lib.b200_paged_attention_workspace_bytes.argtypes = [ctypes.POINTER(PagedAttnArgs)]
    lib.b200_paged_attention_num_launches.argtypes = [ctypes.POINTER(PagedAttnArgs)]
    lib.b200_paged_attention.argtypes = [ctypes.POINTER(PagedAttnArgs), ctypes.c_void_p]
    lib.b200_paged_attention_static_scratch.argtypes = [ctypes.POINTER(PagedAttnArgs), ctypes.c_void_p]
    vp, i32, i64, f32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float
    lib.b200_reshape_and_cache.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, i64, i64, i64, i32, vp]
    lib.b200_copy_blocks.argtypes = [vp, vp, vp, vp, i32, i64, i64, vp]
    lib.b200_qlinear.argtypes = [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i64, i64, vp]
    lib.b200_qlinear_prepacked.argtypes = lib.b200_qlinear.argtypes
    lib.b200_qlinear_experts.argtypes = [vp, vp, vp, vp, vp, vp, ctypes.POINTER(ctypes.c_int32), i32, i32, i32, i32, i32,
                                         i32, i32, i32, i64, i64, vp]
    lib.b200_qlinear_prepack.argtypes = [vp, vp, i32, i32, i32, i32, vp]
    lib.b200_dequantize_rows.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp]
    lib.b200_qstream_blob_bytes.restype = ctypes.c_int64
    lib.b200_qstream_blob_bytes.argtypes = [i32, i32, i32, i32, i32]
    lib.b200_qstream_workspace_bytes.restype = ctypes.c_int64
    lib.b200_qstream_prepack.argtypes = [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp]
    lib.b200_qstream_unpack.argtypes = [vp, vp, i32, i32, i32, i32, i32, vp]
    lib.b200_qstream_linear.argtypes = [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i64, i64, vp, i64, vp]
    lib.b200_cast_bf16_to_f16.argtypes = [vp, vp, i32, i32, i64, vp]
    lib.b200_set_qstream_grid.argtypes = [i32]
    lib.b200_set_qstream_trace.argtypes = [vp]
    lib.b200_rms_norm.argtypes = [vp, vp, vp, vp, i32, i32, i64, i64, i64, f32, i32, i32, vp]
    lib.b200_rope.argtypes = [vp, vp, vp, i32, i32, i32, i32, i32, f32, i32, i64, i64, i32, vp]
    lib.b200_norm_rope_cache.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, f32, i32,
                                         f32, i64, i64, i64, i32, i64, i64, i64, i32, vp]
    lib.b200_silu_mul.argtypes = [vp, vp, i32, i32, i64, i64, i32, vp]
    lib.b200_argmax.argtypes = [vp, vp, i32, i32, i64, i32, vp]
    lib.b200_add_inplace.argtypes = [vp, vp, i32, i32, i64, i64, i32, vp]
    lib.b200_advance_decode.argtypes = [vp, vp, vp, vp, vp, vp, i32, i32, i32, vp]
    if lib.b200_partition_size() != PARTITION_SIZE:
        raise RuntimeError("PARTITION_SIZE mismatch between constants.py and the library")
    return lib


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t: torch.Tensor | None) -> int | None:
    return None if t is None else t.data_ptr()


def _require_cuda(name: str, t: torch.Tensor) -> None:
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name} must be a torch.Tensor, got {type(t).__name__}")
    if not t.is_cuda:
        # reference: PagedAttentionPrimitive::eval_cpu throws (paged_ops.cpp:740-743)
        raise RuntimeError(f"{name} must live on a CUDA device; this path has no CPU implementation")


def _dtype_code(name: str, t: torch.Tensor) -> int:
    try:
        return _DTYPE_CODE[t.dtype]
    except KeyError:
        # reference: dtype_to_metal throws for unsupported dtypes (paged_ops.cpp:185-196)
        raise RuntimeError(f"{name}: unsupported dtype {t.dtype}; expected float32/float16/bfloat16") from None


class _Ops:
    """The operator object ``get_ops()`` returns (reference: module ``_paged_ops``)."""

    PARTITION_SIZE = PARTITION_SIZE

    def __init__(self, lib: ctypes.CDLL) -> None:
        self._lib = lib
        self._workspaces: dict[int, list[torch.Tensor]] = {}
        self._qstream_ws: dict[int, torch.Tensor] = {}
        self.launches = 0  # kernels launched through this object (bench.py's gpu_launches)
        self._qlinear_kernel = 0

    # -- error plumbing -----------------------------------------------------
    def _check(self, rc: int) -> None:
        if rc == 0:
            return
        msg = (self._lib.b200_last_error() or b"").decode()
        if rc in (_ERR_INVALID_ARGUMENT, _ERR_UNSUPPORTED):
            raise ValueError(msg)  # std::invalid_argument -> ValueError in the reference
        raise RuntimeError(msg)

    def _workspace(self, device: torch.device, nbytes: int) -> torch.Tensor | None:
        if nbytes <= 0:
            return None
        idx = device.index if device.index is not None else torch.cuda.current_device()
        pool = self._workspaces.setdefault(idx, [])
        if pool and pool[-1].numel() >= nbytes:
            return pool[-1]
        # keep older buffers alive: captured CUDA graphs may still point at them
        pool.append(torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=device))
        return pool[-1]

    # -- device queries -----------------------------------------------------
    def min_decode_grid(self) -> int:
        n = self._lib.b200_min_decode_grid()
        if n < 0:
            self._check(n)
        return n

    def sm_count(self) -> int:
        n = self._lib.b200_sm_count()
        if n < 0:
            self._check(n)
        return n

    # -- reshape_and_cache (paged_ops.cpp:1718-1749) --------------------------
    def reshape_and_cache(self, key, value, key_cache, value_cache, slot_mapping):
        """Scatter ``key``/``value`` rows into the paged caches by slot.

        Updates the caches in place and returns them; callers rebind as with the
        reference (``sdpa.py:638``).  Negative slots are skipped.
        """
        for n, t in (("key", key), ("value", value), ("key_cache", key_cache),
                     ("value_cache", value_cache), ("slot_mapping", slot_mapping)):
            _require_cuda(n, t)
        if key.dim() != 3 or value.shape != key.shape:
            raise ValueError(f"key/value must both be [T, H_kv, D], got {tuple(key.shape)} / {tuple(value.shape)}")
        if key_cache.dim() != 4 or value_cache.shape != key_cache.shape:
            raise ValueError("key_cache/value_cache must both be [num_blocks, block_size, H_kv, D]")
        if key_cache.shape[2:] != key.shape[1:]:
            raise ValueError(f"cache row shape {tuple(key_cache.shape[2:])} != key row shape {tuple(key.shape[1:])}")
        if key.dtype != key_cache.dtype or value.dtype != value_cache.dtype or key.dtype != value.dtype:
            raise ValueError("key/value must already be cast to the cache dtype")
        if slot_mapping.dtype != torch.int64 or slot_mapping.dim() != 1 or slot_mapping.shape[0] != key.shape[0]:
            raise ValueError("slot_mapping must be int64 [T]")
        if key_cache.stride()[2:] != (key_cache.shape[3], 1) or value_cache.stride() != key_cache.stride():
            raise ValueError("cache rows [H_kv, D] must be contiguous and K/V strides equal")
        key = key.contiguous()
        value = value.contiguous()
        slot_mapping = slot_mapping.contiguous()
        nb, bs, h, d = key_cache.shape
        rc = self._lib.b200_reshape_and_cache(
            key.data_ptr(), value.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(),
            slot_mapping.data_ptr(), key.shape[0], h, d, bs, nb,
            key_cache.stride(0), key_cache.stride(1), _dtype_code("key", key), _stream())
        self._check(rc)
        self.launches += 1
        return key_cache, value_cache

    def copy_blocks(self, key_cache, value_cache, src, dst) -> None:
        """``cache[dst] = cache[src]`` for K and V (kv_cache.py:327-354)."""
        _require_cuda("key_cache", key_cache)
        if not key_cache.is_contiguous() or not value_cache.is_contiguous():
            raise ValueError("copy_blocks needs contiguous caches")
        src = src.to(device=key_cache.device, dtype=torch.int32).contiguous()
        dst = dst.to(device=key_cache.device, dtype=torch.int32).contiguous()
        block_bytes = key_cache.stride(0) * key_cache.element_size()
        rc = self._lib.b200_copy_blocks(key_cache.data_ptr(), value_cache.data_ptr(), src.data_ptr(),
                                        dst.data_ptr(), src.numel(), block_bytes, key_cache.shape[0], _stream())
        self._check(rc)
        self.launches += 1

    # -- paged attention (paged_ops.cpp:1790-1854) ----------------------------
    def paged_attention_primitive(
        self,
        query: torch.Tensor,
        key_cache: torch.Tensor,
        value_cache: torch.Tensor,
        num_kv_heads: int,
        scale: float,
        softcap: float,
        block_tables: torch.Tensor,
        seq_lens: torch.Tensor,
        cu_seqlens_q: torch.Tensor,
        block_size: int,
        max_seq_len: int,
        sliding_window: int,
        out: torch.Tensor,
        key_scale_cache: Any = None,
        value_scale_cache: Any = None,
        key_zero_cache: Any = None,
        v_centroids: Any = None,
        use_turboquant: bool = False,
        quant_type: str = "",
        v_bits: int = 3,
        window_seqlen_q: int = 1,
        sinks: torch.Tensor | None = None,
        *,
        _force_partitions: int = 0,
        _kernel: int = 0,
    ) -> None:
        """Unified paged varlen attention; fills ``out`` in place.

        Same positional/keyword surface as the reference binding.  TurboQuant KV
        compression is out of scope for this build and raises.
        """
        if use_turboquant or key_scale_cache is not None or value_scale_cache is not None:
            if sinks is not None:
                raise ValueError(
                    "attention sinks are not supported with TurboQuant quantized KV; "
                    "pass sinks=None or disable TurboQuant for this layer")
            raise NotImplementedError("TurboQuant KV compression is out of scope of the B200 decode path")
        for n, t in (("query", query), ("key_cache", key_cache), ("value_cache", value_cache),
                     ("block_tables", block_tables), ("seq_lens", seq_lens),
                     ("cu_seqlens_q", cu_seqlens_q), ("out", out)):
            _require_cuda(n, t)
        if query.dim() != 3:
            raise ValueError(f"query must be [T, H_q, D], got {tuple(query.shape)}")
        if sinks is not None:
            # paged_ops.cpp:812-840
            _require_cuda("sinks", sinks)
            if sinks.dim() != 1:
                raise ValueError(f"sinks must be 1-D with one entry per query head, got ndim={sinks.dim()}")
            if sinks.shape[0] != query.shape[1]:
                raise ValueError(
                    f"sinks must have one entry per query head ({query.shape[1]}), got {sinks.shape[0]}")
            if sinks.dtype != torch.float32:
                raise ValueError(
                    "sinks must be float32; the kernel reads them as device float and "
                    "folds them into a float32 softmax accumulator")
            sinks = sinks.contiguous()
        if window_seqlen_q < 1:
            raise ValueError(
                f"window_seqlen_q must be >= 1 (1 = per-token decode), got {window_seqlen_q}; "
                "a non-positive hint would silently select non-window routing")
        if window_seqlen_q > 1:
            # paged_ops.cpp:855-899: the hint must equal the longest segment of a structurally valid cu_seqlens_q
            # (an understated one leaves window rows unwritten); only the spec-verify path passes it and its
            # cu_seqlens_q is a small host-built array, so reading it back is cheap
            if query.shape[2] > PA_WINDOW_MAX_HEAD:
                raise ValueError(
                    f"window_seqlen_q={window_seqlen_q} requires head_size <= {PA_WINDOW_MAX_HEAD}, got "
                    f"{query.shape[2]}; wider heads must keep the expanded per-token verify layout "
                    "(window-mode register state scales with rows * head_size)")
            if cu_seqlens_q.dtype != torch.int32:
                raise ValueError("cu_seqlens_q must be int32 for window-mode validation")
            # the segment check reads cu_seqlens_q back; a stream under CUDA-graph capture cannot, and the eager
            # warm-up call that precedes every capture has already checked the same metadata
            capturing = torch.cuda.is_current_stream_capturing()
            cu = [0, query.shape[0]] if capturing else cu_seqlens_q.tolist()
            window_check = not capturing
            nseg = len(cu) - 1
            if nseg < 1 or cu[0] != 0 or cu[nseg] != query.shape[0]:
                raise ValueError(
                    f"cu_seqlens_q must start at 0 and end at the query row count ({query.shape[0]}), got "
                    f"{nseg} segment(s)")
            longest = 0
            for i in range(nseg):
                seg = cu[i + 1] - cu[i]
                if seg < 1:
                    raise ValueError(
                        f"cu_seqlens_q segment {i} is empty ({cu[i]} -> {cu[i + 1]}); every window-mode segment "
                        "needs at least one query row")
                longest = max(longest, seg)
            if window_check and longest != window_seqlen_q:
                raise ValueError(
                    f"window_seqlen_q={window_seqlen_q} does not match the longest cu_seqlens_q segment "
                    f"({longest}); an understated value leaves window rows unwritten")
        if key_cache.dim() != 4 or value_cache.shape != key_cache.shape:
            raise ValueError("key_cache/value_cache must both be [num_blocks, block_size, H_kv, D]")
        if query.dtype != key_cache.dtype or value_cache.dtype != key_cache.dtype:
            raise ValueError("query and the KV caches must share one dtype (sdpa.py:542-544 casts q/k/v to it)")
        f16_out = out.dtype == torch.float16 and query.dtype == torch.bfloat16  # B200 extension: feeds the fp16-operand o_proj
        if out.shape != query.shape or (out.dtype != query.dtype and not f16_out):
            raise ValueError("out must have the shape and dtype of query (paged_ops.cpp:918-920)")
        t_total, h_q, d = query.shape
        nb, cbs, h_kv, d_c = key_cache.shape
        if cbs != block_size:
            raise ValueError(f"block_size={block_size} does not match the cache's block dim {cbs}")
        if d_c != d or h_kv != num_kv_heads:
            raise ValueError(f"cache [.., {h_kv}, {d_c}] does not match num_kv_heads={num_kv_heads}, head_dim={d}")
        if block_tables.dtype != torch.int32 or block_tables.dim() != 2:
            raise ValueError("block_tables must be int32 [num_seqs, max_blocks]")
        if seq_lens.dtype != torch.int32 or cu_seqlens_q.dtype != torch.int32:
            raise ValueError("seq_lens and cu_seqlens_q must be int32")
        num_seqs = cu_seqlens_q.shape[0] - 1
        if seq_lens.shape[0] != num_seqs or block_tables.shape[0] != num_seqs:
            raise ValueError("seq_lens / block_tables rows must match len(cu_seqlens_q) - 1")
        if query.stride(2) != 1 or query.stride(1) != d or out.stride(2) != 1 or out.stride(1) != d:
            raise ValueError("query/out heads must be contiguous ([T, H_q, D] with unit head stride)")
        if key_cache.stride(3) != 1 or value_cache.stride() != key_cache.stride():
            raise ValueError("cache head_dim must be contiguous and K/V strides equal")
        block_tables = block_tables.contiguous()
        seq_lens = seq_lens.contiguous()
        cu_seqlens_q = cu_seqlens_q.contiguous()

        a = PagedAttnArgs()
        a.out, a.query = out.data_ptr(), query.data_ptr()
        a.key_cache, a.value_cache = key_cache.data_ptr(), value_cache.data_ptr()
        a.block_tables, a.seq_lens = block_tables.data_ptr(), seq_lens.data_ptr()
        a.cu_seqlens_q = cu_seqlens_q.data_ptr()
        a.sinks = _ptr(sinks)
        a.num_q_tokens, a.num_seqs = t_total, num_seqs
        a.num_q_heads, a.num_kv_heads, a.head_dim = h_q, num_kv_heads, d
        a.block_size = block_size
        a.max_blocks_per_seq = max(1, block_tables.shape[1])
        a.max_seq_len = int(max_seq_len)
        a.sliding_window = int(sliding_window)
        a.dtype = _dtype_code("query", query)
        a.force_partitions = int(_force_partitions)
        a.kernel = int(_kernel)
        a.q_stride, a.out_stride = query.stride(0), out.stride(0)
        a.kv_block_stride, a.kv_token_stride, a.kv_head_stride = key_cache.stride()[:3]
        a.num_blocks = nb
        a.scale, a.softcap = float(scale), float(softcap)
        a.out_dtype = _dtype_code("out", out)
        a.window_seqlen_q = int(window_seqlen_q)
        need = self._lib.b200_paged_attention_workspace_bytes(ctypes.byref(a))
        ws = self._workspace(query.device, need)
        a.workspace = _ptr(ws)
        a.workspace_bytes = 0 if ws is None else ws.numel()
        self._check(self._lib.b200_paged_attention(ctypes.byref(a), _stream()))
        self.launches += self._lib.b200_paged_attention_num_launches(ctypes.byref(a))  # from the call's own plan

    # -- MLX-affine quantized linears (mlx_native.py:212-239) -------------------
    def quantized_matmul(self, x, w, scales, biases, transpose: bool = True, group_size: int = 64,
                         bits: int = 4, bias=None, out=None, prepacked: bool = False):
        """``mx.quantized_matmul(x, w, scales, biases, transpose=True, group_size, bits)``.

        ``prepacked=True``: ``w`` is in the streaming layout made by ``qlinear_prepack``."""
        if not transpose:
            raise NotImplementedError("only transpose=True (x @ W.T) is on the hot path")
        for n, t in (("x", x), ("w", w), ("scales", scales), ("biases", biases)):
            _require_cuda(n, t)
        if w.dtype not in (torch.uint32, torch.int32) or w.dim() != 2:
            raise ValueError("w must be uint32 [N, K*bits/32]")
        n_out, words = w.shape
        k = words * 32 // bits
        if x.shape[-1] != k:
            raise ValueError(f"x last dim {x.shape[-1]} != K={k} implied by the packed weight")
        if scales.shape != (n_out, k // group_size) or biases.shape != scales.shape:
            raise ValueError(f"scales/biases must be [{n_out}, {k // group_size}], got {tuple(scales.shape)}")
        if scales.dtype != biases.dtype:
            raise ValueError("scales and biases must share a dtype")
        lead = x.shape[:-1]
        x2 = x.reshape(-1, k)
        if x2.stride(1) != 1:
            x2 = x2.contiguous()
        m = x2.shape[0]
        # result dtype: x's, except an fp16 x into bf16 scales (the tensor-core kernels' operand type) gives bf16
        odt = torch.bfloat16 if (x.dtype == torch.float16 and scales.dtype == torch.bfloat16) else x.dtype
        if out is None:
            out = torch.empty((m, n_out), dtype=odt, device=x.device)
        y2 = out.reshape(m, n_out)
        if bias is not None:
            bias = bias.to(dtype=odt).contiguous()
        n_launch = 1
        if (x2.dtype == torch.bfloat16 and scales.dtype == torch.bfloat16 and k % 64 == 0
                and x2.data_ptr() % 16 == 0 and x2.stride(0) % 8 == 0 and self._qlinear_kernel != 1):
            x2 = self.cast_f16(x2)  # the tcgen05 kernel multiplies fp16 x fp16 (see b200_qlinear)
            self.launches -= 1
            n_launch = 2
        fn = self._lib.b200_qlinear_prepacked if prepacked else self._lib.b200_qlinear
        rc = fn(
            x2.data_ptr(), w.contiguous().data_ptr(), scales.contiguous().data_ptr(),
            biases.contiguous().data_ptr(), _ptr(bias), y2.data_ptr(), m, n_out, k, group_size, bits,
            _dtype_code("x", x2), _dtype_code("out", out), _dtype_code("scales", scales), x2.stride(0), y2.stride(0),
            _stream())
        self._check(rc)
        self.launches += n_launch
        return out.reshape(*lead, n_out)

    def quantized_matmul_experts(self, x, w, scales, biases, counts, group_size: int = 64, bits: int = 4, bias=None):
        """``mx.gather_qmm`` over rows grouped by expert: ``x`` ``[P, K]`` (bf16 / fp16), stacked ``w`` ``[E, N,
        K*bits/32]``, ``scales`` / ``biases`` ``[E, N, K/g]``, ``bias`` ``[E, N]`` or None, ``counts`` host ints [E]
        (rows per expert, in order).  One C call; one kernel launch per expert that has rows."""
        for n, t in (("x", x), ("w", w), ("scales", scales), ("biases", biases)):
            _require_cuda(n, t)
        e, n_out, words = w.shape
        k = words * 32 // bits
        if x.dim() != 2 or x.shape[1] != k or x.stride(1) != 1:
            raise ValueError(f"x must be [P, {k}] with unit inner stride, got {tuple(x.shape)}")
        if len(counts) != e or sum(counts) != x.shape[0]:
            raise ValueError(f"counts must hold {e} row counts adding up to {x.shape[0]}")
        if scales.shape != (e, n_out, k // group_size) or biases.shape != scales.shape or scales.dtype != biases.dtype:
            raise ValueError(f"scales/biases must be [{e}, {n_out}, {k // group_size}] of one dtype")
        odt = torch.bfloat16 if (x.dtype == torch.float16 and scales.dtype == torch.bfloat16) else x.dtype
        out = torch.empty((x.shape[0], n_out), dtype=odt, device=x.device)
        if bias is not None:
            bias = bias.to(dtype=odt).contiguous()
        c_arr = (ctypes.c_int32 * e)(*[int(c) for c in counts])
        rc = self._lib.b200_qlinear_experts(
            x.data_ptr(), w.contiguous().data_ptr(), scales.contiguous().data_ptr(), biases.contiguous().data_ptr(),
            _ptr(bias), out.data_ptr(), c_arr, e, n_out, k, group_size, bits, _dtype_code("x", x), _dtype_code("out", out),
            _dtype_code("scales", scales), x.stride(0), out.stride(0), _stream())
        self._check(rc)
        self.launches += sum(1 for c in counts if c)
        return out

    def qlinear_prepack(self, w, bits: int = 4, inverse: bool = False, out=None):
        """Load-time repack of an MLX-layout packed weight into the streaming layout (same shape)."""
        _require_cuda("w", w)
        if w.dtype not in (torch.uint32, torch.int32) or w.dim() != 2:
            raise ValueError("w must be uint32 [N, K*bits/32]")
        w = w.contiguous()
        if out is None:
            out = torch.empty_like(w)
        n_out, words = w.shape
        self._check(self._lib.b200_qlinear_prepack(w.data_ptr(), out.data_ptr(), n_out, words * 32 // bits, bits,
                                                   int(bool(inverse)), _stream()))
        self.launches += 1
        return out

    # -- streaming-layout linears: persistent stream-K kernel (csrc/qgemm_stream.cu) -----------
    def qstream_prepack(self, w, scales, biases, group_size: int = 64, bits: int = 4, swiglu: bool = False):
        """MLX-affine triple -> the streaming blob (uint8 tensor) read by ``qstream_linear``.
        Load-time; raises ``ValueError`` when a scale/bias does not fit fp16."""
        for n, t in (("w", w), ("scales", scales), ("biases", biases)):
            _require_cuda(n, t)
        if w.dtype not in (torch.uint32, torch.int32) or w.dim() != 2:
            raise ValueError("w must be uint32 [N, K*bits/32]")
        n_out, words = w.shape
        k = words * 32 // bits
        if scales.shape != (n_out, k // group_size) or biases.shape != scales.shape or scales.dtype != biases.dtype:
            raise ValueError(f"scales/biases must be [{n_out}, {k // group_size}] of one dtype")
        nbytes = self._lib.b200_qstream_blob_bytes(n_out, k, bits, group_size, int(swiglu))
        if nbytes <= 0:
            raise ValueError(f"qstream_prepack: unsupported shape N={n_out} K={k} bits={bits} group={group_size}")
        blob = torch.empty((nbytes,), dtype=torch.uint8, device=w.device)
        self._check(self._lib.b200_qstream_prepack(
            w.contiguous().data_ptr(), scales.contiguous().data_ptr(), biases.contiguous().data_ptr(),
            blob.data_ptr(), n_out, k, bits, group_size, _dtype_code("scales", scales), int(swiglu), _stream()))
        self.launches += 1
        return blob

    def qstream_unpack(self, blob, n_out: int, k: int, group_size: int = 64, bits: int = 4, swiglu: bool = False):
        """The MLX-layout packed words back from a streaming blob (bit-exact)."""
        _require_cuda("blob", blob)
        out = torch.empty((n_out, k * bits // 32), dtype=torch.int32, device=blob.device)
        self._check(self._lib.b200_qstream_unpack(blob.data_ptr(), out.data_ptr(), n_out, k, bits, group_size,
                                                  int(swiglu), _stream()))
        self.launches += 1
        return out

    def _qstream_workspace(self, device: torch.device) -> torch.Tensor:
        idx = device.index if device.index is not None else torch.cuda.current_device()
        ws = self._qstream_ws.get(idx)
        if ws is None:
            with torch.cuda.device(idx):
                ws = torch.zeros((self._lib.b200_qstream_workspace_bytes(),), dtype=torch.uint8, device=device)
            self._qstream_ws[idx] = ws
        return ws

    def qstream_linear(self, x, blob, n_out: int, k: int, group_size: int = 64, bits: int = 4, bias=None,
                       swiglu: bool = False, out=None, out_dtype: torch.dtype | None = None,
                       round_dtype: torch.dtype | None = None):
        """``x @ W'^T (+ bias)`` over a streaming blob; ``swiglu=True`` returns ``silu(gate) * up``
        (``[.., n_out / 2]``) from a blob packed with ``swiglu=True``.  ``x`` may be fp16 while the result
        is bf16 (``out_dtype``): the kernel's operands are fp16, a bf16 ``x`` is staged as fp16 first.
        ``out_dtype=float16, round_dtype=bfloat16``: the bf16 result stored as fp16 (feeds the next call)."""
        _require_cuda("x", x)
        _require_cuda("blob", blob)
        if x.shape[-1] != k:
            raise ValueError(f"x last dim {x.shape[-1]} != K={k}")
        lead = x.shape[:-1]
        x2 = x.reshape(-1, k)
        if x2.stride(1) != 1:
            x2 = x2.contiguous()
        m = x2.shape[0]
        cols = n_out // 2 if swiglu else n_out
        if out is None:
            out = torch.empty((m, cols), dtype=out_dtype or x.dtype, device=x.device)
        y2 = out.reshape(m, cols)
        rdt = round_dtype or out.dtype
        store_f16 = rdt == torch.bfloat16 and out.dtype == torch.float16
        if rdt != out.dtype and not store_f16:
            raise ValueError(f"unsupported rounding/storage dtypes {rdt} / {out.dtype}")
        if bias is not None:
            bias = bias.to(dtype=rdt).contiguous()
        if x2.dtype == torch.bfloat16 and m * k * 2 > (8 << 20):
            x2 = self.cast_f16(x2)  # larger than the in-workspace staging region
        ws = self._qstream_workspace(x.device)
        rc = self._lib.b200_qstream_linear(
            x2.data_ptr(), blob.data_ptr(), _ptr(bias), y2.data_ptr(), m, n_out, k, group_size, bits,
            _dtype_code("x", x2), _DTYPE_CODE[rdt], int(swiglu) | (2 if store_f16 else 0), x2.stride(0), y2.stride(0),
            ws.data_ptr(), ws.numel(), _stream())
        self._check(rc)
        self.launches += 2 if x2.dtype == torch.bfloat16 else 1
        return out.reshape(*lead, cols)

    def cast_f16(self, x):
        """bf16 ``[rows, cols]`` -> contiguous fp16, saturating at +-65504 (the stream-K kernel's operand type)."""
        _require_cuda("x", x)
        if x.dtype != torch.bfloat16 or x.dim() != 2 or x.stride(1) != 1:
            raise ValueError("cast_f16 takes a 2-D bf16 tensor with unit inner stride")
        y = torch.empty(x.shape, dtype=torch.float16, device=x.device)
        self._check(self._lib.b200_cast_bf16_to_f16(x.data_ptr(), y.data_ptr(), x.shape[0], x.shape[1], x.stride(0),
                                                    _stream()))
        self.launches += 1
        return y

    def set_qstream_grid(self, grid: int = 0) -> None:
        """Force the persistent grid of the stream-K kernel (0 = heuristic); sweep/test knob."""
        self._check(self._lib.b200_set_qstream_grid(int(grid)))

    def set_qstream_trace(self, buf=None) -> None:
        self._check(self._lib.b200_set_qstream_trace(None if buf is None else buf.data_ptr()))

    def set_pdl(self, enabled: bool = True) -> None:
        """Programmatic dependent launch for every kernel of the library (default on)."""
        self._check(self._lib.b200_set_pdl(int(bool(enabled))))

    def set_qlinear_trace(self, buf=None) -> None:
        """Per-CTA %globaltimer stamps of the tcgen05 qlinear into an int64 CUDA tensor (None = off)."""
        self._lib.b200_set_qlinear_trace.argtypes = [ctypes.c_void_p]
        self._check(self._lib.b200_set_qlinear_trace(None if buf is None else buf.data_ptr()))

    def set_qlinear_variant(self, variant: int = 0) -> None:
        """Ring configuration of the M <= 64 regime (sweep knob, see the header)."""
        self._check(self._lib.b200_set_qlinear_variant(int(variant)))

    def set_qlinear_kernel(self, kernel: int) -> None:
        """0 = auto, 1 = mma.sync kernel, 2 = tcgen05 kernel (raises if the shape does not qualify)."""
        self._check(self._lib.b200_set_qlinear_kernel(int(kernel)))
        self._qlinear_kernel = int(kernel)

    def set_qlinear_tuning(self, ksplit: int = 0, dequant_warps: int = 0) -> None:
        """Sweep knob of the tcgen05 kernel's M <= 64 regime (0 = heuristic)."""
        self._check(self._lib.b200_set_qlinear_tuning(int(ksplit), int(dequant_warps)))

    def dequantize(self, w, scales, biases, group_size: int = 64, bits: int = 4, ids=None,
                   dtype: torch.dtype | None = None):
        """``mx.dequantize(w[ids], scales[ids], biases[ids], group_size, bits)``."""
        for n, t in (("w", w), ("scales", scales), ("biases", biases)):
            _require_cuda(n, t)
        n_out, words = w.shape
        k = words * 32 // bits
        dtype = dtype or scales.dtype
        shape = None
        if ids is not None:
            shape = tuple(ids.shape)
            ids = ids.reshape(-1).to(dtype=torch.int32).contiguous()
            rows = ids.numel()
        else:
            rows = n_out
        out = torch.empty((rows, k), dtype=dtype, device=w.device)
        rc = self._lib.b200_dequantize_rows(
            w.contiguous().data_ptr(), scales.contiguous().data_ptr(), biases.contiguous().data_ptr(),
            _ptr(ids), out.data_ptr(), rows, n_out, k, group_size, bits,
            _dtype_code("scales", scales), _dtype_code("out", out), _stream())
        self._check(rc)
        self.launches += 1
        return out if shape is None else out.reshape(*shape, k)

    # -- neighbours of the hot path (include/b200_decoder_ops.h) ----------------
    def rms_norm(self, x, weight, eps: float, residual=None, out=None, out_dtype: torch.dtype | None = None):
        """``mx.fast.rms_norm``; with ``residual`` does ``residual += x`` first (in place).
        ``out_dtype=torch.float16`` on bf16 input stores the bf16-rounded result as fp16 (the stream-K GEMM's
        operand type)."""
        _require_cuda("x", x)
        dim = x.shape[-1]
        x2 = x.reshape(-1, dim)
        if out is None:
            out = torch.empty(x2.shape, dtype=out_dtype or x.dtype, device=x.device)
        o2 = out.reshape(-1, dim)
        r2 = None if residual is None else residual.reshape(-1, dim)
        rc = self._lib.b200_rms_norm(
            x2.data_ptr(), _ptr(r2), _ptr(weight), o2.data_ptr(), x2.shape[0], dim, x2.stride(0),
            0 if r2 is None else r2.stride(0), o2.stride(0), float(eps), _dtype_code("x", x), _dtype_code("out", out),
            _stream())
        self._check(rc)
        self.launches += 1
        return out.reshape(x.shape)

    def rope(self, q, k, positions, rot_dim: int, base: float, traditional: bool = False):
        """In-place rotary embedding of q [T,Hq,D] and k [T,Hk,D] at int32 positions [T]."""
        _require_cuda("q", q)
        if positions.dtype != torch.int32:
            raise ValueError("positions must be int32")
        t, hq, d = q.shape
        hk = 0 if k is None else k.shape[1]
        rc = self._lib.b200_rope(q.data_ptr(), _ptr(k), positions.data_ptr(), t, hq, hk, d, rot_dim,
                                 float(base), int(traditional), q.stride(0),
                                 0 if k is None else k.stride(0), _dtype_code("q", q), _stream())
        self._check(rc)
        self.launches += 1

    def norm_rope_cache(self, q, k, v, key_cache, value_cache, slot_mapping, positions=None,
                        rot_dim: int = 0, base: float = 10000.0, traditional: bool = False,
                        q_norm_weight=None, k_norm_weight=None, eps: float = 1e-6):
        """Fused per-head RMSNorm -> RoPE -> cache write (sdpa.py:348-368,:638 in one launch).

        ``q [T,Hq,D]``, ``k``/``v [T,Hk,D]`` may be strided views of one qkv buffer (head rows
        contiguous); q and k are updated in place, rotated k and v land in the caches.
        """
        for n, t in (("q", q), ("k", k), ("v", v), ("key_cache", key_cache), ("value_cache", value_cache),
                     ("slot_mapping", slot_mapping)):
            _require_cuda(n, t)
        if q.dim() != 3 or k.dim() != 3 or v.shape != k.shape:
            raise ValueError("q must be [T, Hq, D]; k and v must both be [T, Hk, D]")
        t, hq, d = q.shape
        hk = k.shape[1]
        for n, x in (("q", q), ("k", k), ("v", v)):
            if x.stride(2) != 1 or x.stride(1) != d:
                raise ValueError(f"{n}: heads must be contiguous ([T, H, D] with unit head stride)")
        if key_cache.dim() != 4 or value_cache.shape != key_cache.shape or key_cache.shape[2:] != (hk, d):
            raise ValueError("key_cache/value_cache must both be [num_blocks, block_size, Hk, D]")
        if not (q.dtype == k.dtype == v.dtype == key_cache.dtype == value_cache.dtype):
            raise ValueError("q/k/v must already be in the cache dtype")
        if key_cache.stride()[2:] != (d, 1) or value_cache.stride() != key_cache.stride():
            raise ValueError("cache rows [Hk, D] must be contiguous and K/V strides equal")
        if slot_mapping.dtype != torch.int64 or slot_mapping.shape != (t,):
            raise ValueError("slot_mapping must be int64 [T]")
        if rot_dim and (positions is None or positions.dtype != torch.int32):
            raise ValueError("positions must be int32 [T] when rot_dim > 0")
        nb, bs = key_cache.shape[:2]
        rc = self._lib.b200_norm_rope_cache(
            q.data_ptr(), k.data_ptr(), v.data_ptr(), _ptr(q_norm_weight), _ptr(k_norm_weight),
            _ptr(positions), slot_mapping.contiguous().data_ptr(), key_cache.data_ptr(),
            value_cache.data_ptr(), t, hq, hk, d, int(rot_dim), float(base), int(traditional), float(eps),
            q.stride(0), k.stride(0), v.stride(0), bs, nb, key_cache.stride(0), key_cache.stride(1),
            _dtype_code("q", q), _stream())
        self._check(rc)
        self.launches += 1
        return key_cache, value_cache

    def silu_mul(self, gate_up, out=None):
        _require_cuda("gate_up", gate_up)
        inter = gate_up.shape[-1] // 2
        g2 = gate_up.reshape(-1, 2 * inter)
        if out is None:
            out = torch.empty((g2.shape[0], inter), dtype=gate_up.dtype, device=gate_up.device)
        rc = self._lib.b200_silu_mul(g2.data_ptr(), out.data_ptr(), g2.shape[0], inter, g2.stride(0),
                                     out.stride(0), _dtype_code("gate_up", gate_up), _stream())
        self._check(rc)
        self.launches += 1
        return out

    def argmax(self, logits, out=None):
        """Greedy token ids (reference: mlx_greedy_tokens, v1/sampling_batch.py:342-348)."""
        _require_cuda("logits", logits)
        l2 = logits.reshape(-1, logits.shape[-1])
        if out is None:
            out = torch.empty((l2.shape[0],), dtype=torch.int32, device=logits.device)
        rc = self._lib.b200_argmax(l2.data_ptr(), out.data_ptr(), l2.shape[0], l2.shape[1], l2.stride(0),
                                   _dtype_code("logits", logits), _stream())
        self._check(rc)
        self.launches += 1
        return out

    def add_(self, y, x):
        _require_cuda("y", y)
        dim = y.shape[-1]
        y2, x2 = y.reshape(-1, dim), x.reshape(-1, dim)
        rc = self._lib.b200_add_inplace(y2.data_ptr(), x2.data_ptr(), y2.shape[0], dim, y2.stride(0),
                                        x2.stride(0), _dtype_code("y", y), _stream())
        self._check(rc)
        self.launches += 1
        return y


def _advance_decode(self, input_ids, positions, seq_lens, slot_mapping, block_tables, sampled, block_size):
    """Feed sampled tokens back and move every decode request one slot on (device side)."""
    _require_cuda("input_ids", input_ids)
    if block_tables.stride(1) != 1:
        raise ValueError("block_tables rows must be contiguous")
    rc = self._lib.b200_advance_decode(
        input_ids.data_ptr(), positions.data_ptr(), seq_lens.data_ptr(), slot_mapping.data_ptr(),
        block_tables.data_ptr(), sampled.data_ptr(), input_ids.shape[0], block_tables.stride(0),
        int(block_size), _stream())
    self._check(rc)
    self.launches += 1


_Ops.advance_decode = _advance_decode


def get_ops() -> _Ops:
    """Load (once) and return the native operator object; raises if unavailable."""
    global _ops
    if _ops is None:
        _ops = _Ops(load_library())
        logger.info("Native paged-attention CUDA kernels loaded (%s)", library_path())
    return _ops


def warm_up_kernels() -> None:
    """Front-load library loading at startup (reference: metal/__init__.py:304-327)."""
    try:
        get_ops()
    except Exception as e:  # noqa: BLE001
        raise RuntimeError(f"Failed to load the sm_100a paged-attention kernels: {e}") from e


__all__ = ["get_ops", "warm_up_kernels", "constants", "PARTITION_SIZE", "PagedAttnArgs", "EXPORTED_SYMBOLS"]
