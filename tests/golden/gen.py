"""Portable, seeded input generators shared by the golden-vector script and the tests.

Everything here is integer arithmetic followed by one correctly rounded conversion, so the same
bits come out on any machine / numpy version (no libm, no torch RNG).  The golden file stores a
sha256 of every generated input so that a drift would be noticed.
"""
from __future__ import annotations

import hashlib
import math

import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def _splitmix64(z: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        z = (z + np.uint64(0x9E3779B97F4A7C15)) & _M64
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
        return z ^ (z >> np.uint64(31))


def bits24(seed: int, n: int, stream: int = 0) -> np.ndarray:
    """n integers in [0, 2^24), counter based."""
    with np.errstate(over="ignore"):
        ctr = np.arange(n, dtype=np.uint64) + np.uint64((seed * 0x1000003 + stream * 0x10001) & 0xFFFFFFFFFFFFFFFF) * np.uint64(0x2545F4914F6CDD1D)
    return (_splitmix64(ctr) >> np.uint64(40)).astype(np.int64)


def uniform(seed: int, shape, lo: float, hi: float) -> np.ndarray:
    n = int(np.prod(shape))
    u = bits24(seed, n).astype(np.float64) / float(1 << 24)
    return (lo + (hi - lo) * u).astype(np.float32).reshape(shape)


def normalish(seed: int, shape) -> np.ndarray:
    """Irwin-Hall(12) - 6: mean 0, variance 1, exact integer sums -> float32."""
    n = int(np.prod(shape))
    acc = np.zeros(n, dtype=np.int64)
    for s in range(12):
        acc += bits24(seed, n, stream=s + 1)
    return (acc.astype(np.float64) / float(1 << 24) - 6.0).astype(np.float32).reshape(shape)


def codebooks(seed: int, K: int, C: int, D: int):
    """kaiming-uniform-like U(-sqrt(6/D), sqrt(6/D)) codebooks (core_vq.py:36-39), list of [C, D]."""
    bound = math.sqrt(6.0 / D)
    return [uniform(seed * 101 + k, (C, D), -bound, bound) for k in range(K)]


def randint(seed: int, shape, high: int) -> np.ndarray:
    n = int(np.prod(shape))
    return (bits24(seed, n) % high).astype(np.int64).reshape(shape)


def sha(a: np.ndarray) -> str:
    """sha256 of the bytes; NaNs are canonicalised first (their payload/sign differs between
    x86 and the GPU and carries no meaning)."""
    a = np.ascontiguousarray(a)
    if a.dtype.kind == "f" and np.isnan(a).any():
        a = a.copy()
        a[np.isnan(a)] = np.float32(np.nan)
    return hashlib.sha256(a.tobytes()).hexdigest()[:16]


# --------------------------------------------------------------------------------------------
# Cases.  name -> dict(D, C, K (layers), n_q (active), B, T, seed, variant)
# Shapes come from BASELINE.json configs and from the reference's own call sites
# (tests/quantization/test_vq.py:16, models/builders.py:277, modules/conditioners.py:897-929,1037).
# --------------------------------------------------------------------------------------------
CASES = {
    "c1_musicgen_1x1500": dict(D=128, C=2048, K=4, B=1, T=1500, seed=11),
    "c2_slice_4x1500": dict(D=128, C=2048, K=4, B=4, T=1500, seed=12),
    "c3_deep_k32": dict(D=128, C=1024, K=32, B=2, T=150, seed=13),
    "c5_k8": dict(D=128, C=2048, K=8, B=2, T=256, seed=14),
    "test_vq_shape": dict(D=16, C=8, K=8, B=1, T=2048, seed=15),
    "debug_model": dict(D=32, C=400, K=4, B=8, T=128, seed=16),
    "style_cond_d512_noncontig": dict(D=512, C=1024, K=6, B=2, T=50, seed=17, variant="noncontig"),
    "joint_embed_t1": dict(D=64, C=512, K=12, B=33, T=1, seed=18),
    "d256_default": dict(D=256, C=1024, K=8, B=2, T=77, seed=19),
    "ragged_tail": dict(D=128, C=2048, K=4, B=3, T=131, seed=20),
    "n_q_subset": dict(D=128, C=1024, K=8, n_q=3, B=2, T=100, seed=21),
    "trained_like": dict(D=128, C=2048, K=4, B=2, T=400, seed=22, variant="trained_like"),
    "dup_rows": dict(D=128, C=256, K=2, B=1, T=64, seed=23, variant="dup_rows"),
    "zero_codebook": dict(D=32, C=64, K=2, B=1, T=40, seed=24, variant="zero_codebook"),
    "zero_input": dict(D=128, C=512, K=3, B=1, T=33, seed=25, variant="zero_input"),
    "exact_rows_and_midpoints": dict(D=128, C=512, K=2, B=1, T=96, seed=26, variant="exact_rows"),
    "nonfinite_frames": dict(D=128, C=512, K=3, B=1, T=64, seed=27, variant="nonfinite"),
    "large_scale_inputs": dict(D=128, C=1024, K=4, B=1, T=200, seed=28, variant="large_scale"),
    "tiny_scale_inputs": dict(D=128, C=1024, K=4, B=1, T=200, seed=29, variant="tiny_scale"),
    "empty_T0": dict(D=128, C=256, K=2, B=2, T=0, seed=30),
}


def build_case(name: str):
    """Returns dict(x [B,D,T] fp32 (maybe a non-contiguous view), embed list of [C,D], n_q)."""
    c = dict(CASES[name])
    D, C, K, B, T, seed = c["D"], c["C"], c["K"], c["B"], c["T"], c["seed"]
    variant = c.get("variant", "")
    embed = codebooks(seed, K, C, D)
    x = normalish(seed + 1000, (B, D, T))
    if variant == "noncontig":
        # StyleConditioner passes out.transpose(1, 2) of a [B, T, D] tensor (conditioners.py:956-958)
        x = normalish(seed + 1000, (B, T, D)).transpose(0, 2, 1)
    elif variant == "trained_like":
        # x = sum_k E_k[random code] + small noise: residuals shrink stage by stage
        x = 0.05 * normalish(seed + 1000, (B, D, T))
        for k in range(K):
            scale = np.float32(0.5 ** k)
            embed[k] = (embed[k] * scale).astype(np.float32)
            idx = randint(seed + 2000 + k, (B, T), C)
            x = (x + embed[k][idx].transpose(0, 2, 1)).astype(np.float32)
    elif variant == "dup_rows":
        for k in range(K):
            embed[k][7] = embed[k][5]
            embed[k][40] = embed[k][5]
            embed[k][200] = embed[k][199]
        # make a third of the frames sit exactly on the duplicated row
        x[0, :, 0:20] = embed[0][5][:, None]
    elif variant == "zero_codebook":
        embed = [np.zeros((C, D), dtype=np.float32) for _ in range(K)]
    elif variant == "zero_input":
        x = np.zeros((B, D, T), dtype=np.float32)
    elif variant == "exact_rows":
        idx = randint(seed + 3, (T,), C)
        x[0] = embed[0][idx].T
        half = T // 2
        j = (idx[half:] + 1) % C
        x[0, :, half:] = (np.float32(0.5) * (embed[0][idx[half:]] + embed[0][j])).T
    elif variant == "nonfinite":
        x[0, 3, 5] = np.nan
        x[0, :, 9] = np.nan
        x[0, 10, 17] = np.inf
        x[0, 11, 23] = -np.inf
        x[0, 0, 31] = np.inf
        x[0, 1, 31] = -np.inf
        embed[1][100, 7] = np.nan
        embed[2][3, 0] = np.inf
    elif variant == "large_scale":
        x = (x * np.float32(3.0e4)).astype(np.float32)
        embed = [(e * np.float32(1.0e4)).astype(np.float32) for e in embed]
    elif variant == "tiny_scale":
        x = (x * np.float32(1.0e-6)).astype(np.float32)
        embed = [(e * np.float32(3.0e-6)).astype(np.float32) for e in embed]
    return dict(x=x, embed=embed, n_q=c.get("n_q", K), cfg=c)
