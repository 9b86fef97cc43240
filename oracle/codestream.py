"""numpy restatement of the code-stream plumbing (TEST INFRASTRUCTURE ONLY; the product never imports oracle/).

Follows audiocraft/models/encodec.py:474-486 (stereo interleave) and
audiocraft/modules/codebooks_patterns.py:131-183, 185-246, 339-356 (delayed patterns: scatter tables, build, revert).
Pinned against the imported reference by tests/test_codestream_oracle.py.
"""
from __future__ import annotations

import numpy as np


def stereo_interleave(c0: np.ndarray, c1: np.ndarray, per_timestep: bool) -> np.ndarray:
    """'c b k t -> b (k c) t' / 'c b k t -> b k (t c)' with c = 2 (encodec.py:476-479)."""
    st = np.stack([c0, c1], axis=0)                      # [c, b, k, t]
    c, b, k, t = st.shape
    if per_timestep:
        return st.transpose(1, 2, 3, 0).reshape(b, k, t * c)
    return st.transpose(1, 2, 0, 3).reshape(b, k * c, t)


def stereo_split(joint: np.ndarray, per_timestep: bool):
    """encodec.py:483-486."""
    if per_timestep:
        b, k, tc = joint.shape
        x = joint.reshape(b, k, tc // 2, 2).transpose(3, 0, 1, 2)
    else:
        b, kc, t = joint.shape
        x = joint.reshape(b, kc // 2, 2, t).transpose(2, 0, 1, 3)
    return x[0], x[1]


def delay_layout(delays, offset: int, timesteps: int):
    """codebooks_patterns.py:339-356 with flatten_first == 0: list of [(t, q), ...] per sequence step."""
    out = [[] for _ in range(offset)]
    for t in range(timesteps + max(delays)):
        out.append([(t - d, q) for q, d in enumerate(delays) if t - d >= 0])
    return out


def delay_build(z: np.ndarray, delays, offset: int, special: int, keep_only_valid_steps: bool):
    """codebooks_patterns.py:131-183 (timesteps of the pattern == z's)."""
    B, K, T = z.shape
    layout = delay_layout(delays, offset, T)
    if keep_only_valid_steps:
        layout = layout[: len(layout) - max(delays)]
    S = len(layout)
    indexes = np.full((K, S), K * T, dtype=np.int64)
    mask = np.zeros((K, S), dtype=bool)
    for s, coords in enumerate(layout):
        for (t, q) in coords:
            if t < T:
                indexes[q, s] = t + q * T
                mask[q, s] = True
    flat = np.concatenate([z.reshape(B, -1), np.full((B, 1), special, dtype=z.dtype)], axis=1)
    return flat[:, indexes.reshape(-1)].reshape(B, K, S), indexes, mask


def delay_revert(seq: np.ndarray, delays, offset: int, special: int, timesteps: int, keep_only_valid_steps: bool):
    """codebooks_patterns.py:185-246."""
    B, K, S = seq.shape
    layout = delay_layout(delays, offset, timesteps)
    if keep_only_valid_steps:
        layout = layout[: len(layout) - max(delays)]
    assert S <= len(layout)
    indexes = np.full((K, timesteps), K * S, dtype=np.int64)
    mask = np.zeros((K, timesteps), dtype=bool)
    for s, coords in enumerate(layout):
        if s < S:
            for (t, q) in coords:
                if t < timesteps:
                    indexes[q, t] = s + q * S
                    mask[q, t] = True
    flat = np.concatenate([seq.reshape(B, -1), np.full((B, 1), special, dtype=seq.dtype)], axis=1)
    return flat[:, indexes.reshape(-1)].reshape(B, K, timesteps), indexes, mask


def pack_bits(codes: np.ndarray, bits: int) -> np.ndarray:
    """n codes -> ceil(n * bits / 8) bytes; code j occupies bits [j * bits, (j + 1) * bits) of the stream, little endian."""
    n = codes.size
    out = np.zeros((n * bits + 7) // 8, dtype=np.uint8)
    v = codes.astype(np.uint64) & np.uint64((1 << bits) - 1)
    for b in range(bits):
        pos = np.arange(n, dtype=np.int64) * bits + b
        bit = ((v >> np.uint64(b)) & np.uint64(1)).astype(np.uint8)
        np.bitwise_or.at(out, pos >> 3, bit << (pos & 7).astype(np.uint8))
    return out
