"""The RVQ of HuggingFace `transformers`' EnCodec on the B200 kernels.

`HFEncodecCompressionModel` (audiocraft/models/encodec.py:323-394) wraps `transformers.EncodecModel`; its
`decode_latent` calls `self.model.quantizer.decode(codes.transpose(0, 1))` (:360-362) and its `encode` / `decode` go
through the same `EncodecResidualVectorQuantizer` (transformers/models/encodec/modeling_encodec.py), which is the
formula and the four-buffer layout of audiocraft's own quantizer under other names:

    transformers                                   audiocraft / this package
    quantizer.layers.{i}.codebook.{inited,         vq.layers.{i}._codebook.{inited,
        cluster_size, embed, embed_avg}                cluster_size, embed, embed_avg}
    encode(embeddings [B, D, T], bandwidth) -> [K, B, T]      encode(x) -> [B, K, T] (a view of that block)
    decode(codes [K, B, T]) -> [B, D, T]                       decode(codes [B, K', T]) -> [B, D, T]

`HFEncodecQuantizer` has the transformers class's surface and state_dict keys and runs on `ResidualVectorQuantizer`;
`patch_hf_model(model)` swaps it into a loaded `EncodecModel` (`facebook/encodec_24khz`, `encodec_32khz`, ...).
`transformers` itself is not imported here: the adapter only needs the module's attributes.
"""
from __future__ import annotations

import math
import typing as tp

import torch
from torch import nn

from .quantization import ResidualVectorQuantizer

_BUFFERS = ("inited", "cluster_size", "embed", "embed_avg")


class HFEncodecQuantizer(nn.Module):
    """Drop-in for `transformers.models.encodec.modeling_encodec.EncodecResidualVectorQuantizer` (eval)."""

    def __init__(self, codebook_size: int, codebook_dim: int, num_quantizers: int, frame_rate: float):
        super().__init__()
        self.codebook_size = codebook_size
        self.frame_rate = frame_rate
        self.num_quantizers = num_quantizers
        self.rvq = ResidualVectorQuantizer(dimension=codebook_dim, n_q=num_quantizers, bins=codebook_size,
                                           kmeans_init=False).eval()

    # ---- state --------------------------------------------------------------------------------------------------
    @staticmethod
    def _to_ours(key: str) -> str:
        # layers.3.codebook.embed -> rvq.vq.layers.3._codebook.embed
        parts = key.split(".")
        assert parts[0] == "layers" and parts[2] == "codebook" and parts[3] in _BUFFERS, key
        return f"rvq.vq.layers.{parts[1]}._codebook.{parts[3]}"

    def state_dict(self, *args, prefix: str = "", **kwargs):  # transformers' key names
        ours = super().state_dict(*args, prefix="", **kwargs)
        out = type(ours)()
        for k, v in ours.items():
            p = k.split(".")  # rvq.vq.layers.i._codebook.name
            out[f"{prefix}layers.{p[3]}.codebook.{p[5]}"] = v
        return out

    def load_state_dict(self, state_dict, strict: bool = True, **kwargs):
        return super().load_state_dict({self._to_ours(k): v for k, v in state_dict.items()}, strict=strict, **kwargs)

    @classmethod
    def from_hf(cls, hf_quantizer: nn.Module) -> "HFEncodecQuantizer":
        """From a `transformers` EncodecResidualVectorQuantizer (buffers copied, device kept)."""
        first = hf_quantizer.layers[0].codebook.embed
        q = cls(int(hf_quantizer.codebook_size), int(first.shape[1]), int(hf_quantizer.num_quantizers),
                float(hf_quantizer.frame_rate))
        q.load_state_dict({k: v for k, v in hf_quantizer.state_dict().items()})
        return q.to(first.device)

    # ---- the transformers surface -------------------------------------------------------------------------------
    def get_num_quantizers_for_bandwidth(self, bandwidth: tp.Optional[float] = None) -> int:
        bw_per_q = math.log2(self.codebook_size) * self.frame_rate
        num_quantizers = self.num_quantizers
        if bandwidth is not None and bandwidth > 0.0:
            num_quantizers = int(max(1, math.floor(bandwidth * 1000 / bw_per_q)))
        return num_quantizers

    @torch.no_grad()
    def encode(self, embeddings: torch.Tensor, bandwidth: tp.Optional[float] = None) -> torch.Tensor:
        """[B, D, T] -> codes [K, B, T] (all K stages in one kernel launch)."""
        n_q = min(self.get_num_quantizers_for_bandwidth(bandwidth), self.num_quantizers)
        return self.rvq.vq.encode(embeddings, n_q=n_q)

    @torch.no_grad()
    def decode(self, codes: torch.Tensor) -> torch.Tensor:
        """codes [K', B, T] -> [B, D, T] (one gather-sum launch)."""
        return self.rvq.vq.decode(codes)


def patch_hf_model(model: nn.Module) -> nn.Module:
    """Replaces `model.quantizer` of a loaded `transformers.EncodecModel` by the kernel-backed twin."""
    model.quantizer = HFEncodecQuantizer.from_hf(model.quantizer)
    return model
