"""The two convolutions that touch the quantizer in `EncodecModel`, in the layout the kernels like.

`SEANetEncoder` ends with `StreamableConv1d(mult * n_filters, dimension, last_kernel_size)` (modules/seanet.py:143-148)
and hands `[B, D, T]` with T contiguous to `quantizer.encode` (models/encodec.py:236-237); `SEANetDecoder` starts with
`StreamableConv1d(dimension, mult * n_filters, kernel_size)` (modules/seanet.py:209-213) on the quantizer's output,
which is physically `[B, T, D]` (core_vq.py:399-405 sums permuted views).  The encode kernel reads frame-major
`[B, T, D]` latents without its in-kernel transposition (about 5 % faster, profiles/r2_notes.md), and the decoder conv
can consume the `[B, T, D]` result as it is.  `EdgeConv1d` is the same convolution as `StreamableConv1d` (stride 1,
same asymmetric / causal padding, modules/conv.py:187-203) evaluated on channels-last memory:

    out = EdgeConv1d.from_reference(encoder.model[-1])(h)      # [B, D, T] view of a [B, T, D] block: strides (T*D, 1, D)
    codes = quantizer.encode(out)                               # no permute, no copy, no in-kernel transposition
    z = quantizer.decode(codes)                                 # [B, D, T] view of [B, T, D]
    y = EdgeConv1d.from_reference(decoder.model[0])(z)          # consumes that layout directly

The convolution itself is a library call (cuDNN through torch, channels-last conv2d with a height of 1): it is not part
of the hot path, only its output / input layout is.
"""
from __future__ import annotations

import math

import torch
from torch import nn
from torch.nn import functional as F


class EdgeConv1d(nn.Module):
    """Stride-1 `StreamableConv1d` (modules/conv.py:167-203) whose input and output live in channels-last memory."""

    def __init__(self, weight: torch.Tensor, bias: torch.Tensor | None, dilation: int = 1, causal: bool = False,
                 pad_mode: str = "reflect"):
        super().__init__()
        assert weight.dim() == 3
        self.weight = nn.Parameter(weight.detach().clone()[:, :, None, :].contiguous(memory_format=torch.channels_last),
                                   requires_grad=False)
        self.bias = None if bias is None else nn.Parameter(bias.detach().clone(), requires_grad=False)
        self.dilation = int(dilation)
        self.causal = causal
        self.pad_mode = pad_mode

    @classmethod
    def from_reference(cls, streamable_conv: nn.Module) -> "EdgeConv1d":
        """From a reference `StreamableConv1d` (any `norm`: the effective weight is read through the parametrization)."""
        conv = streamable_conv.conv.conv
        assert conv.stride[0] == 1 and conv.groups == 1, "the edge convolutions of SEANet have stride 1 and no groups"
        return cls(conv.weight, conv.bias, dilation=conv.dilation[0], causal=streamable_conv.causal,
                   pad_mode=streamable_conv.pad_mode)

    def _pad_frames(self, x_btd: torch.Tensor, left: int, right: int) -> torch.Tensor:
        """pad1d (modules/conv.py:72-89) along the frame axis of a [B, T, C] block."""
        if left == 0 and right == 0:
            return x_btd
        if self.pad_mode != "reflect":
            return F.pad(x_btd, (0, 0, left, right), mode=self.pad_mode)
        T = x_btd.shape[1]
        extra = 0
        if T <= max(left, right):          # tiny inputs: zero-extend on the right first, cut it off afterwards
            extra = max(left, right) - T + 1
            x_btd = F.pad(x_btd, (0, 0, 0, extra))
        T2 = x_btd.shape[1]
        parts = []
        if left:
            parts.append(x_btd[:, 1:left + 1].flip(1))
        parts.append(x_btd)
        if right:
            parts.append(x_btd[:, T2 - 1 - right:T2 - 1].flip(1))
        out = torch.cat(parts, dim=1)
        return out[:, :out.shape[1] - extra] if extra else out

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """x [B, C_in, T] (any strides; frame-major memory is consumed without a copy) -> [B, C_out, T] whose memory
        is [B, T, C_out]."""
        B, Cin, T = x.shape
        k = (self.weight.shape[-1] - 1) * self.dilation + 1
        total = k - 1                                   # kernel_size - stride, stride == 1: no extra padding needed
        left, right = (total, 0) if self.causal else (total - total // 2, total // 2)
        x_btd = x.transpose(1, 2)
        if not x_btd.is_contiguous():
            x_btd = x_btd.contiguous()
        x_btd = self._pad_frames(x_btd, left, right)    # [B, T + total, C_in]
        x4 = x_btd.permute(0, 2, 1)[:, :, None, :]      # [B, C_in, 1, T + total] with channels-last strides
        y4 = F.conv2d(x4, self.weight, self.bias, dilation=(1, self.dilation))
        y4 = y4.contiguous(memory_format=torch.channels_last)   # a no-op when cuDNN already answered channels-last
        return y4[:, :, 0, :]                           # [B, C_out, T], strides (T * C_out, 1, C_out)
