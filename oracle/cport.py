"""ctypes front-end of the oracle's C restatement (``oracle/csrc/oracle_cpu.c``).
TEST / BASELINE INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

Arrays are NumPy; 16-bit tensors are bf16 bit patterns in ``uint16`` (use
``oracle._numeric.bf16_bits`` / ``bf16_from_bits``).
"""

from __future__ import annotations

import ctypes

import numpy as np

from . import build as _build

_vp = ctypes.c_void_p


def _p(a):
    return None if a is None else a.ctypes.data_as(_vp)


def _c(a, dt):
    a = np.ascontiguousarray(a, dtype=dt)
    return a


def num_threads() -> int:
    return int(_build.load().oracle_num_threads())


def set_num_threads(n: int) -> int:
    """Use ``n`` OpenMP threads from now on (torchrun exports OMP_NUM_THREADS=1; the CPU arm of
    bench.py must not inherit that).  Returns the team size in effect."""
    _build.load().oracle_set_num_threads(ctypes.c_int(int(n)))
    return num_threads()


def reshape_and_cache(key, value, key_cache, value_cache, slot_mapping):
    lib = _build.load()
    assert key_cache.flags.c_contiguous and value_cache.flags.c_contiguous
    t = key.shape[0]
    row = int(np.prod(key.shape[1:]))
    k, v, s = _c(key, np.uint16), _c(value, np.uint16), _c(slot_mapping, np.int64)
    lib.oracle_reshape_and_cache(_p(k), _p(v), _p(key_cache), _p(value_cache), _p(s), ctypes.c_int(t),
                                 ctypes.c_int(row))
    return key_cache, value_cache


def paged_attention(q, key_cache, value_cache, block_tables, seq_lens, cu_seqlens_q, scale,
                    sliding_window=-1, softcap=0.0, sinks=None, out=None):
    lib = _build.load()
    t, hq, d = q.shape
    _, bs, hkv, _ = key_cache.shape
    q = _c(q, np.uint16)
    bt = _c(block_tables, np.int32)
    sl, cu = _c(seq_lens, np.int32), _c(cu_seqlens_q, np.int32)
    sk = None if sinks is None else _c(sinks, np.float32)
    if out is None:
        out = np.zeros_like(q)
    lib.oracle_paged_attention(
        _p(out), _p(q), _p(key_cache), _p(value_cache), _p(bt), _p(sl), _p(cu), _p(sk),
        ctypes.c_int(len(sl)), ctypes.c_int(hq), ctypes.c_int(hkv), ctypes.c_int(d), ctypes.c_int(bs),
        ctypes.c_int(bt.shape[1] if bt.ndim == 2 and bt.shape[1] else 1), ctypes.c_int(sliding_window),
        ctypes.c_float(scale), ctypes.c_float(softcap))
    return out


def dequantize(w, scales, biases, group, bits, ids=None):
    lib = _build.load()
    n, words = w.shape
    k = words * 32 // bits
    ids_c = None if ids is None else _c(ids, np.int32)
    rows = n if ids is None else len(ids_c)
    out = np.empty((rows, k), np.uint16)
    lib.oracle_dequantize(_p(out), _p(_c(w, np.uint32)), _p(_c(scales, np.uint16)), _p(_c(biases, np.uint16)),
                          _p(ids_c), ctypes.c_int(rows), ctypes.c_int(k), ctypes.c_int(group), ctypes.c_int(bits))
    return out


def qmatmul(x, w, scales, biases, group, bits, out=None):
    lib = _build.load()
    m, k = x.shape
    n = w.shape[0]
    if out is None:
        out = np.empty((m, n), np.uint16)
    lib.oracle_qmatmul(_p(out), _p(_c(x, np.uint16)), _p(_c(w, np.uint32)), _p(_c(scales, np.uint16)),
                       _p(_c(biases, np.uint16)), ctypes.c_int(m), ctypes.c_int(n), ctypes.c_int(k),
                       ctypes.c_int(group), ctypes.c_int(bits))
    return out


def rms_norm(x, weight, eps, residual=None):
    lib = _build.load()
    rows, dim = x.shape
    out = np.empty_like(x)
    lib.oracle_rms_norm(_p(out), _p(_c(x, np.uint16)), _p(residual), _p(None if weight is None else _c(weight, np.uint16)),
                        ctypes.c_int(rows), ctypes.c_int(dim), ctypes.c_float(eps))
    return out


def rope_(x, positions, base):
    lib = _build.load()
    t, h, d = x.shape
    assert x.flags.c_contiguous and x.dtype == np.uint16
    lib.oracle_rope(_p(x), _p(_c(positions, np.int32)), ctypes.c_int(t), ctypes.c_int(h), ctypes.c_int(d),
                    ctypes.c_float(base))
    return x


def silu_mul(gate_up):
    lib = _build.load()
    rows, two = gate_up.shape
    out = np.empty((rows, two // 2), np.uint16)
    lib.oracle_silu_mul(_p(out), _p(_c(gate_up, np.uint16)), ctypes.c_int(rows), ctypes.c_int(two // 2))
    return out


def argmax(logits):
    lib = _build.load()
    rows, vocab = logits.shape
    out = np.empty(rows, np.int32)
    lib.oracle_argmax(_p(out), _p(_c(logits, np.uint16)), ctypes.c_int(rows), ctypes.c_int(vocab))
    return out
