"""Module tree of the residual vector quantizer, state_dict-compatible with the reference.

The reference (audiocraft/quantization/core_vq.py) nests
    ResidualVectorQuantization.layers[i] : VectorQuantization
        ._codebook : EuclideanCodebook with buffers inited[1], cluster_size[C], embed[C, D], embed_avg[C, D]
and checkpoints are loaded strictly by those names (models/loaders.py:176).  The same tree is kept
here, but the arithmetic is not spread over it: the eval-mode work of all stages is done by one
fused kernel call per encode / forward and one per decode (include/rvq_b200.h), issued from
`ResidualVectorQuantization`.  The per-layer classes keep the reference's per-layer methods for
callers that use them directly; they run the same kernels on a one-stage pack.
"""
from __future__ import annotations

import math
import typing as tp

import torch
from torch import nn

from .. import kernels


def _kaiming_uniform_codebook(bins: int, dim: int) -> torch.Tensor:
    # uniform_init (core_vq.py:36-39): kaiming_uniform_ with fan_in = dim  ->  U(-sqrt(6/dim), sqrt(6/dim))
    t = torch.empty(bins, dim)
    nn.init.kaiming_uniform_(t)
    return t


class CodebookPack:
    """Private, derived device copy of a list of `embed` buffers (fp32 block, ||c||^2, fp16 operand
    copies; see rvq_pack_codebooks).  Rebuilt whenever an embed buffer is replaced (`.to()`,
    `load_state_dict` on a fresh module) or written in place (version counter)."""

    def __init__(self) -> None:
        self._key: tp.Optional[tuple] = None
        self._pack: tp.Optional[torch.Tensor] = None
        self._status: tp.Optional[torch.Tensor] = None

    def invalidate(self) -> None:
        self._key = None
        self._pack = None

    def get(self, embeds: tp.Sequence[torch.Tensor]) -> torch.Tensor:
        key = tuple((e.data_ptr(), e._version, e.device) for e in embeds)
        if key != self._key or self._pack is None:
            self._pack = kernels.pack_codebooks(embeds)
            self._key = key
        return self._pack

    def status(self, device: torch.device) -> torch.Tensor:
        if self._status is None or self._status.device != device:
            self._status = torch.zeros(2, dtype=torch.int64, device=device)
        return self._status


def _raise_if_bad_codes(status: torch.Tensor) -> None:
    flag, where = status.tolist()  # device -> host read: synchronises the stream
    if flag:
        status.zero_()
        raise IndexError(f"index out of range in self (flat code index {where - 1})")


class EuclideanCodebook(nn.Module):
    """One codebook and its statistics (core_vq.py:86-127).  Owns the four checkpointed buffers."""

    def __init__(self, dim: int, codebook_size: int, kmeans_init: bool = False, kmeans_iters: int = 10,
                 decay: float = 0.8, epsilon: float = 1e-5, threshold_ema_dead_code: int = 2):
        super().__init__()
        self.decay = decay
        self.codebook_size = codebook_size
        self.kmeans_iters = kmeans_iters
        self.epsilon = epsilon
        self.threshold_ema_dead_code = threshold_ema_dead_code
        embed = torch.zeros(codebook_size, dim) if kmeans_init else _kaiming_uniform_codebook(codebook_size, dim)
        self.register_buffer("inited", torch.Tensor([not kmeans_init]))
        self.register_buffer("cluster_size", torch.zeros(codebook_size))
        self.register_buffer("embed", embed)
        self.register_buffer("embed_avg", embed.clone())
        self._pack = CodebookPack()

    @property
    def dim(self) -> int:
        return self.embed.shape[1]

    # ---- per-layer eval API of the reference (core_vq.py:160-193), channels-last input [..., D]
    def encode(self, x: torch.Tensor) -> torch.Tensor:
        lead = x.shape[:-1]
        flat = x.reshape(1, -1, x.shape[-1]).transpose(1, 2)  # [1, D, N] view, frames are columns
        codes, _ = kernels.encode(flat, self._pack.get([self.embed]), 1, self.codebook_size, 1)
        return codes.view(*lead)

    def decode(self, embed_ind: torch.Tensor) -> torch.Tensor:
        lead = embed_ind.shape
        status = self._pack.status(embed_ind.device)
        out = kernels.decode(embed_ind.reshape(1, 1, -1), self._pack.get([self.embed]), 1, self.codebook_size,
                             self.dim, status)
        _raise_if_bad_codes(status)
        return out.view(*lead, self.dim)


class VectorQuantization(nn.Module):
    """One RVQ stage (core_vq.py:222-349).  RVQ never sets `codebook_dim`, so the in/out projections
    of the reference are Identity (core_vq.py:262-266) and are not instantiated here."""

    def __init__(self, dim: int, codebook_size: int, codebook_dim: tp.Optional[int] = None, decay: float = 0.8,
                 epsilon: float = 1e-5, kmeans_init: bool = False, kmeans_iters: int = 10,
                 threshold_ema_dead_code: int = 2, channels_last: bool = False, commitment_weight: float = 1.,
                 orthogonal_reg_weight: float = 0.0, orthogonal_reg_active_codes_only: bool = False,
                 orthogonal_reg_max_codes: tp.Optional[int] = None):
        super().__init__()
        if codebook_dim is not None and codebook_dim != dim:
            raise NotImplementedError("projected codebooks (codebook_dim != dim) are outside the RVQ path")
        self.project_in = nn.Identity()
        self.project_out = nn.Identity()
        self.epsilon = epsilon
        self.commitment_weight = commitment_weight
        self.orthogonal_reg_weight = orthogonal_reg_weight
        self.orthogonal_reg_active_codes_only = orthogonal_reg_active_codes_only
        self.orthogonal_reg_max_codes = orthogonal_reg_max_codes
        self._codebook = EuclideanCodebook(dim=dim, codebook_size=codebook_size, kmeans_init=kmeans_init,
                                           kmeans_iters=kmeans_iters, decay=decay, epsilon=epsilon,
                                           threshold_ema_dead_code=threshold_ema_dead_code)
        self.codebook_size = codebook_size
        self.channels_last = channels_last

    @property
    def codebook(self) -> torch.Tensor:
        return self._codebook.embed

    @property
    def inited(self) -> torch.Tensor:
        return self._codebook.inited

    def encode(self, x: torch.Tensor) -> torch.Tensor:
        """[B, D, T] (or [B, T, D] when channels_last) -> codes [B, T] (core_vq.py:301-305)."""
        if not self.channels_last:
            x = x.transpose(1, 2)
        return self._codebook.encode(x)

    def decode(self, embed_ind: torch.Tensor) -> torch.Tensor:
        """codes [B, T] -> [B, D, T] view of a [B, T, D] tensor (core_vq.py:307-311)."""
        q = self._codebook.decode(embed_ind)
        return q if self.channels_last else q.transpose(1, 2)


class ResidualVectorQuantization(nn.Module):
    """The K-stage residual chain (core_vq.py:352-405), fused."""

    def __init__(self, *, num_quantizers: int, **kwargs):
        super().__init__()
        self.layers = nn.ModuleList([VectorQuantization(**kwargs) for _ in range(num_quantizers)])
        self._pack = CodebookPack()
        self.encode_algo = "auto"  # "auto" | "exact" | "tc" (include/rvq_b200.h RVQ_ALGO_*)
        self.check_codes = True    # read the decode error word back after every decode (host sync)

    # -- helpers
    def _embeds(self) -> tp.List[torch.Tensor]:
        return [layer._codebook.embed for layer in self.layers]

    def _shape(self) -> tp.Tuple[int, int, int]:
        e = self.layers[0]._codebook.embed
        return len(self.layers), e.shape[0], e.shape[1]

    def invalidate_cache(self) -> None:
        """Call after writing to an `embed` buffer through `.data` (version counters miss that)."""
        self._pack.invalidate()

    def check_errors(self) -> None:
        """Raises IndexError if a decode since the last check saw an out-of-range code."""
        if self._pack._status is not None:
            _raise_if_bad_codes(self._pack._status)

    # -- eval path
    def encode(self, x: torch.Tensor, n_q: tp.Optional[int] = None) -> torch.Tensor:
        """[B, D, T] fp32 -> codes [K, B, T] int64 (core_vq.py:387-397)."""
        Ktot, C, _ = self._shape()
        K = n_q or Ktot
        codes, _ = kernels.encode(x, self._pack.get(self._embeds()), Ktot, C, min(K, Ktot), algo=self.encode_algo)
        return codes

    def decode(self, q_indices: torch.Tensor) -> torch.Tensor:
        """codes [K', B, T] -> [B, D, T] fp32, physically [B, T, D] (core_vq.py:399-405)."""
        Ktot, C, D = self._shape()
        if q_indices.shape[0] > Ktot:
            raise IndexError(f"index {Ktot} is out of range")  # self.layers[i] in the reference
        if q_indices.shape[0] == 0:
            return torch.tensor(0.0, device=q_indices.device)
        status = self._pack.status(q_indices.device)
        out = kernels.decode(q_indices.transpose(0, 1), self._pack.get(self._embeds()), Ktot, C, D, status)
        if self.check_codes:
            _raise_if_bad_codes(status)
        return out.transpose(1, 2)

    def forward(self, x: torch.Tensor, n_q: tp.Optional[int] = None):
        """Eval branch of core_vq.py:363-385: (quantized [B, D, T], codes [K, B, T], losses [K, 1])."""
        if self.training:
            raise NotImplementedError(
                "training-mode RVQ (k-means init, EMA codebook updates, commitment loss) is not part of the "
                "B200 path yet; call .eval() first")
        Ktot, C, _ = self._shape()
        K = min(n_q or Ktot, Ktot)
        codes, quant = kernels.encode(x, self._pack.get(self._embeds()), Ktot, C, K, want_quantized=True,
                                      algo=self.encode_algo)
        losses = torch.zeros((K, 1), device=x.device)
        return quant.transpose(1, 2), codes, losses
