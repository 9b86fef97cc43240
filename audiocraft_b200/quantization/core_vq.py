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


def _broadcast_buffers(module: nn.Module) -> None:
    """Keeps the codebook statistics identical on every worker after a random (re-)initialisation; the
    reference does this through flashy.distrib.broadcast_tensors (core_vq.py:140,158)."""
    if torch.distributed.is_available() and torch.distributed.is_initialized() and \
            torch.distributed.get_world_size() > 1:
        for buf in module.buffers():
            torch.distributed.broadcast(buf, src=0)


def _draw_vectors(samples: torch.Tensor, num: int) -> torch.Tensor:
    """`num` rows of `samples`, without replacement when there are enough (core_vq.py:42-50; same random
    calls, so the same rows under the same seed)."""
    n = samples.shape[0]
    if n >= num:
        pick = torch.randperm(n, device=samples.device)[:num]
    else:
        pick = torch.randint(0, n, (num,), device=samples.device)
    return samples[pick]


def _kmeans(samples: torch.Tensor, num_clusters: int, num_iters: int = 10, chunk: int = 4096):
    """Lloyd iterations from randomly drawn means (core_vq.py:53-75).  The assignment step evaluates the same
    -(sum_d (x_d - m_d)^2) per pair as the reference but over chunks of samples, so that memory stays
    O(chunk * clusters * dim) instead of O(samples * clusters * dim)."""
    dim, dtype = samples.shape[-1], samples.dtype
    means = _draw_vectors(samples, num_clusters)
    bins = torch.zeros(num_clusters, dtype=torch.long, device=samples.device)
    for _ in range(num_iters):
        buckets = torch.empty(samples.shape[0], dtype=torch.long, device=samples.device)
        for lo in range(0, samples.shape[0], chunk):
            part = samples[lo:lo + chunk]
            dists = -((part[:, None, :] - means[None, :, :]) ** 2).sum(dim=-1)
            buckets[lo:lo + chunk] = dists.max(dim=-1).indices
        bins = torch.bincount(buckets, minlength=num_clusters)
        empty = bins == 0
        sums = torch.zeros(num_clusters, dim, dtype=dtype, device=samples.device)
        sums.index_add_(0, buckets, samples)
        new_means = sums / bins.masked_fill(empty, 1)[:, None]
        means = torch.where(empty[:, None], means, new_means)
    return means, bins


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

    # ---- codebook maintenance (training-time statistics, core_vq.py:129-158, 205-217) -------------------
    @torch.no_grad()
    def init_embed_(self, data: torch.Tensor) -> None:
        """k-means initialisation from the first batch seen (not gated by self.training, like the reference)."""
        if bool(self.inited):
            return
        embed, cluster_size = _kmeans(data, self.codebook_size, self.kmeans_iters)
        self.embed.copy_(embed)
        self.embed_avg.copy_(embed)
        self.cluster_size.copy_(cluster_size)
        self.inited.fill_(1.0)
        _broadcast_buffers(self)

    @torch.no_grad()
    def expire_codes_(self, batch_samples: torch.Tensor) -> None:
        """Codes whose EMA usage fell below the threshold are replaced by random vectors of the batch."""
        if self.threshold_ema_dead_code == 0:
            return
        dead = self.cluster_size < self.threshold_ema_dead_code
        if not bool(torch.any(dead)):
            return
        fresh = _draw_vectors(batch_samples.reshape(-1, batch_samples.shape[-1]), self.codebook_size)
        self.embed.copy_(torch.where(dead[:, None], fresh, self.embed))
        _broadcast_buffers(self)

    @torch.no_grad()
    def _ema_update_(self, flat: torch.Tensor, embed_ind: torch.Tensor) -> None:
        """EMA of code usage and of the per-code sum of inputs, then the smoothed re-normalisation
        (core_vq.py:209-217).  The reference forms both statistics through an [N, C] one-hot matrix; here
        they are a histogram and a segmented sum of the same quantities."""
        if flat.is_cuda:
            # one pass of librvq_b200.so: ordered per-code sums (no atomics), EMA and re-normalisation fused
            kernels.ema_update(flat, embed_ind.reshape(-1), self.cluster_size, self.embed_avg, self.embed, self.decay,
                               self.epsilon)
            self._pack.invalidate()  # embed was written through a raw pointer: torch's version counter did not move
            return
        # host tensors: the same statistics with torch ops (the CPU tests compare this host logic with the reference
        # step by step; the search itself has no CPU path)
        C = self.codebook_size
        idx = embed_ind.reshape(-1)
        counts = torch.bincount(idx, minlength=C).to(flat.dtype)
        sums = torch.zeros(C, flat.shape[-1], dtype=flat.dtype, device=flat.device)
        sums.index_add_(0, idx, flat)
        self.cluster_size.mul_(self.decay).add_(counts, alpha=1 - self.decay)
        self.embed_avg.mul_(self.decay).add_(sums, alpha=1 - self.decay)
        n = self.cluster_size.sum()
        smoothed = (self.cluster_size + self.epsilon) / (n + C * self.epsilon) * n
        self.embed.copy_(self.embed_avg / smoothed.unsqueeze(1))

    def _search(self, flat: torch.Tensor) -> torch.Tensor:
        """Nearest code of every row of flat [N, D] -> int64 [N] (the B200 kernels; CUDA only)."""
        codes, _ = kernels.encode(flat.t().unsqueeze(0), self._pack.get([self.embed]), 1, self.codebook_size, 1)
        return codes.view(-1)

    def forward(self, x: torch.Tensor):
        """Per-layer path of the reference (core_vq.py:195-219): channels-last x [..., D] ->
        (quantized [..., D], codes [...]).  Used in training mode and while a codebook is uninitialised."""
        lead = x.shape[:-1]
        flat = x.reshape(-1, x.shape[-1])
        self.init_embed_(flat)
        embed_ind = self._search(flat.detach())
        quantized = torch.nn.functional.embedding(embed_ind, self.embed).view(*lead, -1)
        if self.training:
            self.expire_codes_(flat.detach())
            self._ema_update_(flat.detach(), embed_ind)
        return quantized, embed_ind.view(*lead)

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

    def forward(self, x: torch.Tensor):
        """One stage with its training-time terms (core_vq.py:313-349): (quantized, codes, loss[1])."""
        if not self.channels_last:
            x = x.transpose(1, 2)
        quantized, embed_ind = self._codebook(x)
        if self.training:
            quantized = x + (quantized - x).detach()  # straight-through estimator
        loss = torch.tensor([0.0], device=x.device, requires_grad=self.training)
        if self.training:
            if self.commitment_weight > 0:
                loss = loss + torch.nn.functional.mse_loss(quantized.detach(), x) * self.commitment_weight
            if self.orthogonal_reg_weight > 0:
                codebook = self.codebook
                if self.orthogonal_reg_active_codes_only:
                    codebook = codebook[torch.unique(embed_ind)]
                n_codes = codebook.shape[0]
                if self.orthogonal_reg_max_codes is not None and n_codes > self.orthogonal_reg_max_codes:
                    codebook = codebook[torch.randperm(n_codes, device=x.device)[:self.orthogonal_reg_max_codes]]
                # eq. (2) of arXiv:2112.00384 (core_vq.py:78-84): mean squared off-diagonal cosine similarity
                unit = torch.nn.functional.normalize(codebook, p=2, dim=-1)
                gram = unit @ unit.t()
                n = unit.shape[0]
                ortho = ((gram - torch.eye(n, device=x.device)) ** 2).sum() / (n ** 2)
                loss = loss + ortho * self.orthogonal_reg_weight
        if not self.channels_last:
            quantized = quantized.transpose(1, 2)
        return quantized, embed_ind, loss


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

    def _all_inited(self) -> bool:
        """Host-side copy of the `inited` flags (the reference reads the device flag per layer and call,
        core_vq.py:131); refreshed whenever one of them may have changed."""
        key = tuple((layer._codebook.inited.data_ptr(), layer._codebook.inited._version) for layer in self.layers)
        if getattr(self, "_inited_key", None) != key:
            self._inited_all = all(bool(layer._codebook.inited) for layer in self.layers)
            self._inited_key = key
        return self._inited_all

    def _forward_per_layer(self, x: torch.Tensor, n_q: tp.Optional[int] = None):
        """Stage-by-stage path with the training-time terms (core_vq.py:363-385)."""
        quantized_out: tp.Any = 0.0
        residual = x
        losses, indices = [], []
        n_q = n_q or len(self.layers)
        for layer in self.layers[:n_q]:
            quantized, idx, loss = layer(residual)
            quantized = quantized.detach()
            residual = residual - quantized
            quantized_out = quantized_out + quantized
            indices.append(idx)
            losses.append(loss)
        if self.training:
            # straight-through over the whole chain (facebookresearch/encodec issue 25, core_vq.py:380-382)
            quantized_out = x + (quantized_out - x).detach()
        self._pack.invalidate()
        return quantized_out, torch.stack(indices), torch.stack(losses)

    def invalidate_cache(self) -> None:
        """Call after writing to an `embed` buffer through `.data` (version counters miss that)."""
        self._pack.invalidate()

    def check_errors(self) -> None:
        """Raises IndexError if a decode since the last check saw an out-of-range code."""
        if self._pack._status is not None:
            _raise_if_bad_codes(self._pack._status)

    # -- eval path
    def encode(self, x: torch.Tensor, n_q: tp.Optional[int] = None, wire: tp.Optional[torch.Tensor] = None) -> torch.Tensor:
        """[B, D, T] fp32 -> codes [K, B, T] int64 (core_vq.py:387-397).  `wire` (not in the reference): an int16
        [K, B, T] view that receives the same codes, for the all-gather of a frame-sharded job (sharded.py)."""
        Ktot, C, _ = self._shape()
        K = n_q or Ktot
        codes, _ = kernels.encode(x, self._pack.get(self._embeds()), Ktot, C, min(K, Ktot), algo=self.encode_algo,
                                  wire=wire)
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
        if self.training or not self._all_inited():
            return self._forward_per_layer(x, n_q)
        Ktot, C, _ = self._shape()
        K = min(n_q or Ktot, Ktot)
        codes, quant = kernels.encode(x, self._pack.get(self._embeds()), Ktot, C, K, want_quantized=True,
                                      algo=self.encode_algo)
        losses = torch.zeros((K, 1), device=x.device)
        return quant.transpose(1, 2), codes, losses
