"""Quantizer interface, same surface as the reference's audiocraft/quantization/base.py.

`QuantizedResult` (base.py:18-24), `BaseQuantizer` (base.py:27-65) and the pass-through
`DummyQuantizer` (base.py:68-107) that the reference's EnCodec tests use as a fake backend.
"""
from __future__ import annotations

import typing as tp
from dataclasses import dataclass, field

import torch
from torch import nn


@dataclass
class QuantizedResult:
    x: torch.Tensor                                # quantized latent [B, D, T]
    codes: torch.Tensor                            # [B, K, T] int64
    bandwidth: torch.Tensor                        # kb/s per batch item, 0-dim
    penalty: tp.Optional[torch.Tensor] = None      # commitment loss term (0 in eval)
    metrics: dict = field(default_factory=dict)


class BaseQuantizer(nn.Module):
    """Abstract surface every quantizer offers to EncodecModel (encodec.py:170-259)."""

    def forward(self, x: torch.Tensor, frame_rate: int) -> QuantizedResult:
        raise NotImplementedError()

    def encode(self, x: torch.Tensor) -> torch.Tensor:
        raise NotImplementedError()

    def decode(self, codes: torch.Tensor) -> torch.Tensor:
        raise NotImplementedError()

    @property
    def total_codebooks(self) -> int:
        raise NotImplementedError()

    @property
    def num_codebooks(self) -> int:
        raise NotImplementedError()

    def set_num_codebooks(self, n: int) -> None:
        raise NotImplementedError()


class DummyQuantizer(BaseQuantizer):
    """Identity "quantizer": codes are the latent itself with a singleton codebook axis."""

    def forward(self, x: torch.Tensor, frame_rate: int) -> QuantizedResult:
        codes = x.unsqueeze(1)
        # 32 bits per value, per second, per batch item (base.py:76)
        kbps = codes.numel() * 32 * frame_rate / 1000 / len(x)
        return QuantizedResult(x, codes, torch.tensor(kbps).to(x))

    def encode(self, x: torch.Tensor) -> torch.Tensor:
        return x.unsqueeze(1)

    def decode(self, codes: torch.Tensor) -> torch.Tensor:
        return codes.squeeze(1)

    @property
    def total_codebooks(self) -> int:
        return 1

    @property
    def num_codebooks(self) -> int:
        return self.total_codebooks

    def set_num_codebooks(self, n: int) -> None:
        raise AttributeError("Cannot override the number of codebooks for the dummy quantizer")
