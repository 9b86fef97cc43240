"""ctypes front end of oracle/rvq_oracle.c (TEST INFRASTRUCTURE ONLY).

numpy in, numpy out.  Shapes follow the reference (audiocraft/quantization):
x [B, D, T] fp32 -> codes [B, K, T] int64 (a view of a [K, B, T] block, vq.py:87-96);
codes [B, K', T] -> quantized [B, D, T] fp32 with strides (T*D, 1, D) in elements (vq.py:98-104).
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_build", "librvq_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    """Compile the oracle with the committed Makefile (gcc only)."""
    src = os.path.join(_HERE, "rvq_oracle.c")
    stale = (not os.path.exists(LIB_PATH)) or os.path.getmtime(LIB_PATH) < os.path.getmtime(src)
    if force or stale:
        subprocess.run(["make", "-C", _HERE] + (["-B"] if force else []), check=True,
                       stdout=subprocess.DEVNULL)
    return LIB_PATH


def _load():
    global _lib
    if _lib is None:
        build()
        lib = ctypes.CDLL(LIB_PATH)
        f32p = ctypes.POINTER(ctypes.c_float)
        i64p = ctypes.POINTER(ctypes.c_int64)
        i64, i32 = ctypes.c_int64, ctypes.c_int
        lib.rvq_oracle_cnorm.argtypes = [f32p, i32, i32, i32, f32p]
        lib.rvq_oracle_cnorm.restype = None
        lib.rvq_oracle_encode.argtypes = [f32p, i64, i64, i64, i32, i32, i32, f32p, i32, i32,
                                          i64p, f32p, f32p]
        lib.rvq_oracle_encode.restype = i32
        lib.rvq_oracle_decode.argtypes = [i64p, i64, i64, i64, i32, i32, i32, f32p, i32, i32, f32p]
        lib.rvq_oracle_decode.restype = i64
        lib.rvq_oracle_dot.argtypes = [f32p, f32p, i32]
        lib.rvq_oracle_dot.restype = ctypes.c_float
        lib.rvq_oracle_set_threads.argtypes = [i32]
        lib.rvq_oracle_get_threads.restype = i32
        _lib = lib
    return _lib


def _f32p(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


def _i64p(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_int64))


def set_threads(n: int) -> None:
    _load().rvq_oracle_set_threads(int(n))


def get_threads() -> int:
    return int(_load().rvq_oracle_get_threads())


def _embed_block(embed) -> np.ndarray:
    """list of [C, D] arrays or one [K, C, D] array -> contiguous fp32 [K, C, D]."""
    if isinstance(embed, (list, tuple)):
        embed = np.stack([np.asarray(e, dtype=np.float32) for e in embed])
    embed = np.ascontiguousarray(embed, dtype=np.float32)
    assert embed.ndim == 3
    return embed


def cnorm(embed) -> np.ndarray:
    embed = _embed_block(embed)
    K, C, D = embed.shape
    out = np.empty((K, C), dtype=np.float32)
    _load().rvq_oracle_cnorm(_f32p(embed), K, C, D, _f32p(out))
    return out


def dot(a, b) -> float:
    a = np.ascontiguousarray(a, dtype=np.float32)
    b = np.ascontiguousarray(b, dtype=np.float32)
    return float(_load().rvq_oracle_dot(_f32p(a), _f32p(b), a.shape[0]))


def _encode(x, embed, n_q, want_q, want_scores):
    embed = _embed_block(embed)
    Ktot, C, D = embed.shape
    K = Ktot if n_q is None else int(n_q)
    assert 0 < K <= Ktot
    x = np.asarray(x)
    if x.dtype != np.float32:
        raise TypeError("oracle encode expects float32 input (the reference raises on other dtypes)")
    B, Dx, T = x.shape
    assert Dx == D
    es = x.itemsize
    sb, sd, st = (s // es for s in x.strides)
    kbt = np.empty((K, B, T), dtype=np.int64)
    q = np.empty((B, T, D), dtype=np.float32) if want_q else None
    sc = np.empty((K, B, T), dtype=np.float32) if want_scores else None
    if B * T > 0:
        _load().rvq_oracle_encode(_f32p(x), sb, sd, st, B, D, T, _f32p(embed), K, C, _i64p(kbt),
                                  _f32p(q) if want_q else None,
                                  _f32p(sc) if want_scores else None)
    return kbt, q, sc


def encode(x, embed, n_q=None, return_scores=False):
    """ResidualVectorQuantizer.encode (vq.py:87-96): codes [B, K, T] as a transposed view."""
    kbt, _, sc = _encode(x, embed, n_q, False, return_scores)
    codes = kbt.transpose(1, 0, 2)
    return (codes, sc.transpose(1, 0, 2)) if return_scores else codes


def decode(codes, embed) -> np.ndarray:
    """ResidualVectorQuantizer.decode (vq.py:98-104): [B, K', T] -> [B, D, T] (view of [B, T, D])."""
    embed = _embed_block(embed)
    Ktot, C, D = embed.shape
    codes = np.asarray(codes)
    if codes.dtype != np.int64:
        codes = codes.astype(np.int64)
    B, K, T = codes.shape
    assert K <= Ktot
    es = codes.itemsize
    cb, ck, ct = (s // es for s in codes.strides)
    out = np.empty((B, T, D), dtype=np.float32)
    if K == 0:
        out[...] = 0.0
    elif B * T > 0:
        bad = _load().rvq_oracle_decode(_i64p(codes), cb, ck, ct, B, K, T, _f32p(embed), C, D,
                                        _f32p(out))
        if bad:
            raise IndexError("index out of range in self")
    return out.transpose(0, 2, 1)


def forward(x, embed, n_q=None):
    """Eval-mode ResidualVectorQuantizer.forward (vq.py:76-85): (quantized [B, D, T], codes [B, K, T])."""
    kbt, q, _ = _encode(x, embed, n_q, True, False)
    return q.transpose(0, 2, 1), kbt.transpose(1, 0, 2)
