"""`ResidualVectorQuantizer`: the drop-in boundary (reference: audiocraft/quantization/vq.py:16-116).

Constructed by name from checkpoint configs (`builders.get_quantizer`, models/builders.py:44-53)
with exactly these keyword arguments, called by `EncodecModel` as `quantizer(emb, frame_rate)`,
`.encode(emb)`, `.decode(codes)` (models/encodec.py:212,237,259).
"""
from __future__ import annotations

import math
import typing as tp

import torch

from .base import BaseQuantizer, QuantizedResult
from .core_vq import ResidualVectorQuantization


class ResidualVectorQuantizer(BaseQuantizer):
    """Residual vector quantizer running on the B200 kernels.

    Args (same names, order and defaults as vq.py:35-48):
        dimension: size D of the latent vectors.
        n_q: number of residual stages K.
        q_dropout: training only, draw the number of active stages at random.
        bins: codebook size C.
        decay, kmeans_init, kmeans_iters, threshold_ema_dead_code, orthogonal_reg_*: training-time
            codebook maintenance hyper-parameters, stored and forwarded to the stages.
    """

    def __init__(
        self,
        dimension: int = 256,
        n_q: int = 8,
        q_dropout: bool = False,
        bins: int = 1024,
        decay: float = 0.99,
        kmeans_init: bool = True,
        kmeans_iters: int = 10,
        threshold_ema_dead_code: int = 2,
        orthogonal_reg_weight: float = 0.0,
        orthogonal_reg_active_codes_only: bool = False,
        orthogonal_reg_max_codes: tp.Optional[int] = None,
    ):
        super().__init__()
        self.max_n_q = n_q
        self.n_q = n_q
        self.q_dropout = q_dropout
        self.dimension = dimension
        self.bins = bins
        self.decay = decay
        self.kmeans_init = kmeans_init
        self.kmeans_iters = kmeans_iters
        self.threshold_ema_dead_code = threshold_ema_dead_code
        self.orthogonal_reg_weight = orthogonal_reg_weight
        self.orthogonal_reg_active_codes_only = orthogonal_reg_active_codes_only
        self.orthogonal_reg_max_codes = orthogonal_reg_max_codes
        self.vq = ResidualVectorQuantization(
            dim=dimension, codebook_size=bins, num_quantizers=n_q, decay=decay, kmeans_init=kmeans_init,
            kmeans_iters=kmeans_iters, threshold_ema_dead_code=threshold_ema_dead_code,
            orthogonal_reg_weight=orthogonal_reg_weight,
            orthogonal_reg_active_codes_only=orthogonal_reg_active_codes_only,
            orthogonal_reg_max_codes=orthogonal_reg_max_codes, channels_last=False)
        self._bw_cache: tp.Dict[tuple, torch.Tensor] = {}

    def _bandwidth(self, n_q: int, frame_rate: float, like: torch.Tensor) -> torch.Tensor:
        # kb/s = stages * log2(bins) bits * frames per second / 1000 (vq.py:80,84); cached per device so
        # that repeated calls do not pay a host->device copy each time
        key = (n_q, float(frame_rate), like.device, like.dtype)
        bw = self._bw_cache.get(key)
        if bw is None:
            bw = torch.tensor(n_q * (math.log2(self.bins) * frame_rate / 1000)).to(like)
            self._bw_cache[key] = bw
        return bw

    def forward(self, x: torch.Tensor, frame_rate: int) -> QuantizedResult:
        n_q = self.n_q
        if self.training and self.q_dropout:
            n_q = int(torch.randint(1, self.n_q + 1, (1,)).item())
        quantized, codes, commit_loss = self.vq(x, n_q=n_q)
        # codes come back as [K, B, T]; callers see the [B, K, T] view (vq.py:82)
        return QuantizedResult(quantized, codes.transpose(0, 1), self._bandwidth(n_q, frame_rate, x),
                               penalty=torch.mean(commit_loss))

    def encode(self, x: torch.Tensor) -> torch.Tensor:
        """[B, D, T] fp32 -> codes [B, K, T] int64 using the first `n_q` stages."""
        return self.vq.encode(x, n_q=self.n_q).transpose(0, 1)

    def decode(self, codes: torch.Tensor) -> torch.Tensor:
        """codes [B, K', T] -> quantized latent [B, D, T]."""
        return self.vq.decode(codes.transpose(0, 1))

    @property
    def total_codebooks(self) -> int:
        return self.max_n_q

    @property
    def num_codebooks(self) -> int:
        return self.n_q

    def set_num_codebooks(self, n: int) -> None:
        assert n > 0 and n <= self.max_n_q
        self.n_q = n
