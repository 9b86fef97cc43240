"""What happens to the codes after `ResidualVectorQuantizer.encode`: the host side of
csrc/rvq_codestream.cu, with the reference's names.

  * stereo interleave     `InterleaveStereoCompressionModel.encode` / `get_left_right_codes`
                          (audiocraft/models/encodec.py:474-486)
  * delayed patterns      `DelayedPatternProvider` / `ParallelPatternProvider` / `Pattern.build_pattern_sequence` /
                          `Pattern.revert_pattern_sequence` (audiocraft/modules/codebooks_patterns.py:22-270, 313-368)
  * compact code formats  int16 (`codes.to(torch.int16)`, modules/jasco_conditioners.py:136) and bit-packed
                          (11 bits per code for 2048 bins; not in the reference)

CUDA tensors only, like the rest of the package: every function is one kernel launch of librvq_b200.so.
"""
from __future__ import annotations

import math
import typing as tp
from collections import namedtuple

import torch

from . import kernels

LayoutCoord = namedtuple('LayoutCoord', ['t', 'q'])  # (timestep, codebook index), as in the reference


# ---- stereo ---------------------------------------------------------------------------------------------------
def interleave_stereo_codes(indices_c0: torch.Tensor, indices_c1: torch.Tensor, per_timestep: bool = False) -> torch.Tensor:
    """encodec.py:470-479: `stack([c0, c1])` then 'c b k t -> b (k c) t' (or 'c b k t -> b k (t c)')."""
    return kernels.stereo_interleave(indices_c0, indices_c1, per_timestep)


def get_left_right_codes(codes: torch.Tensor, per_timestep: bool = False) -> tp.Tuple[torch.Tensor, torch.Tensor]:
    """encodec.py:483-486."""
    return kernels.stereo_split(codes, per_timestep)


# ---- compact formats -------------------------------------------------------------------------------------------
def to_int16(codes: torch.Tensor) -> torch.Tensor:
    """[B, K, T] codes (any strides) -> contiguous int16 (the reference's cache format)."""
    return kernels.convert_codes(codes, torch.int16)


def to_int64(codes: torch.Tensor) -> torch.Tensor:
    return kernels.convert_codes(codes, torch.int64)


def bits_for(bins: int) -> int:
    return max(1, math.ceil(math.log2(bins)))


def pack_codes(codes: torch.Tensor, bins: int) -> torch.Tensor:
    """int64 codes in [0, bins) -> uint8 bytes, ceil(log2(bins)) bits per code."""
    return kernels.pack_codes(codes, bits_for(bins))


def unpack_codes(packed: torch.Tensor, shape: tp.Sequence[int], bins: int) -> torch.Tensor:
    return kernels.unpack_codes(packed, shape, bits_for(bins))


# ---- delayed patterns ------------------------------------------------------------------------------------------
class Pattern:
    """The layouts `DelayedPatternProvider.get_pattern` produces (codebooks_patterns.py:339-356 with
    flatten_first == 0): codebook q sits `offset + delays[q]` sequence steps after its timestep.  Same methods and
    return values as the reference's `Pattern` (values, indexes, mask); the scatter tables come out of the same
    kernel launch instead of a Python loop over the layout."""

    def __init__(self, delays: tp.Sequence[int], offset: int, timesteps: int, n_q: int):
        self.delays = list(delays)
        self.offset = int(offset)
        self.timesteps = int(timesteps)
        self.n_q = int(n_q)

    @property
    def max_delay(self) -> int:
        return max(self.delays)

    @property
    def layout(self) -> tp.List[tp.List[LayoutCoord]]:
        """The explicit layout of the reference (for inspection and parity tests; the kernels do not use it)."""
        out: tp.List[tp.List[LayoutCoord]] = [[] for _ in range(self.offset)]
        for t in range(self.timesteps + self.max_delay):
            out.append([LayoutCoord(t - d, q) for q, d in enumerate(self.delays) if t - d >= 0])
        return out

    @property
    def num_sequence_steps(self) -> int:
        return self.offset + self.timesteps + self.max_delay - 1

    @property
    def valid_layout(self):
        return self.layout[: self.offset + self.timesteps]

    def starts_with_special_token(self) -> bool:
        return self.offset > 0

    def build_pattern_sequence(self, z: torch.Tensor, special_token: int, keep_only_valid_steps: bool = False):
        """z [B, K, T] -> (values [B, K, S], indexes [K, S], mask [K, S]) (codebooks_patterns.py:160-183)."""
        B, K, T = z.shape
        assert K == self.n_q, f"invalid number of codebooks for the sequence and the pattern: {K} != {self.n_q}"
        assert T <= self.timesteps, "invalid number of timesteps used to build the sequence from the pattern"
        S = self.offset + self.timesteps + (0 if keep_only_valid_steps else self.max_delay)
        return kernels.delay_pattern(z, self.delays, self.offset, special_token, S, revert=False)

    def revert_pattern_sequence(self, s: torch.Tensor, special_token: int, keep_only_valid_steps: bool = False):
        """s [B, K, S] -> (values [B, K, T], indexes [K, T], mask [K, T]) (codebooks_patterns.py:224-246)."""
        B, K, S = s.shape
        assert K == self.n_q, f"invalid number of codebooks for the sequence and the pattern: {K} != {self.n_q}"
        steps = self.offset + self.timesteps + (0 if keep_only_valid_steps else self.max_delay)
        assert S <= steps, f"sequence to revert is longer than the defined pattern: {S} > {steps}"
        return kernels.delay_pattern(s, self.delays, self.offset, special_token, self.timesteps, revert=True,
                                     valid_steps=steps)


class DelayedPatternProvider:
    """codebooks_patterns.py:313-356.  `flatten_first` is not supported by the kernels (MusicGen does not use it)."""

    def __init__(self, n_q: int, delays: tp.Optional[tp.List[int]] = None, flatten_first: int = 0, empty_initial: int = 0):
        assert n_q > 0
        self.n_q = n_q
        if delays is None:
            delays = list(range(n_q))
        self.delays = list(delays)
        self.flatten_first = flatten_first
        self.empty_initial = empty_initial
        assert len(self.delays) == self.n_q
        assert sorted(self.delays) == self.delays
        if flatten_first:
            raise NotImplementedError("flatten_first != 0 is outside the kernels' delayed layouts")

    def get_pattern(self, timesteps: int) -> Pattern:
        offset = (0 if self.empty_initial < 0 else 1) + max(self.empty_initial, 0)
        return Pattern(self.delays, offset, timesteps, self.n_q)


class ParallelPatternProvider(DelayedPatternProvider):
    """codebooks_patterns.py:359-368: the delayed pattern with no delay."""

    def __init__(self, n_q: int, empty_initial: int = 0):
        super().__init__(n_q, [0] * n_q, empty_initial=empty_initial)


def decode_compact(quantizer, codes16: torch.Tensor) -> torch.Tensor:
    """`quantizer.decode` for int16 codes [B, K, T] (the cache format) without widening them first.  (The reference's
    `decode` rejects int16 indices, so the drop-in method keeps doing that; this is the explicit way in.)"""
    vq = quantizer.vq
    Ktot, C, D = vq._shape()
    status = vq._pack.status(codes16.device)
    out = kernels.decode(codes16, vq._pack.get(vq._embeds()), Ktot, C, D, status, allow_int16=True)
    if vq.check_codes:
        vq.check_errors()
    return out.transpose(1, 2)


def decode_pattern_sequence(quantizer, seq: torch.Tensor, pattern: Pattern) -> torch.Tensor:
    """`quantizer.decode(pattern.revert_pattern_sequence(seq, ...)[0])` in one kernel launch: the gather reads the
    language model's delayed [B, K, S] tokens in place (rvq_decode_delayed_f32).  `seq` must hold every step of the
    pattern's `timesteps` frames (S >= offset + timesteps + max_delay); returns [B, D, T] like `decode`."""
    vq = quantizer.vq
    Ktot, C, D = vq._shape()
    B, K, S = seq.shape
    assert K == pattern.n_q
    late = [pattern.offset + d for d in pattern.delays]
    status = vq._pack.status(seq.device)
    out = kernels.decode(seq, vq._pack.get(vq._embeds()), Ktot, C, D, status, steps_late=late, frames=pattern.timesteps)
    if vq.check_codes:
        vq.check_errors()
    return out.transpose(1, 2)
