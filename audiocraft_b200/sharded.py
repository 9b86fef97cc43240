"""Frame-sharded RVQ over the GPUs of one box (one process per GPU, torch.distributed / NCCL).

The reference's inference path is single-device; this is the multi-GPU design of
BASELINE.json config 5.  RVQ is frame independent -- the K-stage chain is sequential only inside a
frame -- so any split of the B*T frames is valid: no halo, no exchange during compute.

  * codebooks: replicated, one broadcast of the module's buffers from rank 0 at construction
  * compute:   every rank encodes / decodes its contiguous slice (whole batch items when B >= world,
               else a slice of the time axis)
  * exchange:  one all-gather of the codes per call, as int16 on the wire when bins <= 32767
               (4x fewer bytes than int64), widened locally.  Quantized latents are not gathered:
               re-decoding gathered codes locally is cheaper than moving fp32 latents over NVLink.
"""
from __future__ import annotations

import typing as tp

import torch
import torch.distributed as dist


def shard_sizes(n: int, world: int) -> tp.List[int]:
    """Sizes of `world` contiguous shards of n items; the first n % world shards get one more."""
    base, extra = divmod(n, world)
    return [base + (1 if r < extra else 0) for r in range(world)]


def shard_bounds(n: int, world: int, rank: int) -> tp.Tuple[int, int]:
    sizes = shard_sizes(n, world)
    lo = sum(sizes[:rank])
    return lo, lo + sizes[rank]


def shard_axis(B: int, T: int, world: int) -> int:
    """0: split the batch axis (whole items per rank); 2: split the time axis (fewer items than ranks)."""
    return 0 if B >= world else 2


class ShardedRVQ:
    """Wraps a quantizer (any object with the BaseQuantizer surface) for frame-sharded use.

    Every rank constructs it with its own replica of the module; rank 0's buffers win.
    """

    def __init__(self, quantizer, group: tp.Optional[dist.ProcessGroup] = None, wire_int16: tp.Optional[bool] = None):
        self.q = quantizer
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        bins = getattr(quantizer, "bins", 1 << 30)
        self.wire_int16 = (bins <= 32767) if wire_int16 is None else wire_int16
        if self.world > 1 and hasattr(quantizer, "buffers"):
            for buf in quantizer.buffers():
                dist.broadcast(buf, src=0, group=group)

    # ---- local pieces -----------------------------------------------------------------------
    def local_slice(self, x: torch.Tensor) -> torch.Tensor:
        """This rank's part of a global [B, D, T] (or codes [B, K, T]) tensor."""
        B, _, T = x.shape
        axis = shard_axis(B, T, self.world)
        lo, hi = shard_bounds(x.shape[axis], self.world, self.rank)
        return x[lo:hi] if axis == 0 else x[:, :, lo:hi]

    def encode_local(self, x_local: torch.Tensor) -> torch.Tensor:
        return self.q.encode(x_local)

    def decode_local(self, codes_local: torch.Tensor) -> torch.Tensor:
        return self.q.decode(codes_local)

    # ---- exchange ---------------------------------------------------------------------------
    def gather_codes(self, codes_local: torch.Tensor, global_shape: tp.Tuple[int, int, int]) -> torch.Tensor:
        """All-gathers per-rank codes [B_r, K, T_r] into the global [B, K, T] (int64) on every rank."""
        B, K, T = global_shape
        if self.world == 1:
            return codes_local
        axis = shard_axis(B, T, self.world)
        sizes = shard_sizes(B if axis == 0 else T, self.world)
        smax = max(sizes)
        wire = torch.int16 if self.wire_int16 else torch.int64
        # [K, B_r, T_r] is what the encode kernel wrote; pad the sharded axis to the largest shard
        kbt = codes_local.transpose(0, 1)
        pad_shape = (K, smax, T) if axis == 0 else (K, B, smax)
        send = torch.zeros(pad_shape, dtype=wire, device=codes_local.device)
        if axis == 0:
            send[:, :kbt.shape[1]] = kbt
        else:
            send[:, :, :kbt.shape[2]] = kbt
        # concatenation along dim 0 (the layout both NCCL and Gloo accept), viewed as [world, ...] afterwards
        recv = torch.empty((self.world * pad_shape[0],) + pad_shape[1:], dtype=wire, device=codes_local.device)
        # an all-gather only moves bytes: send them as uint8 (neither NCCL nor Gloo has an int16 type)
        dist.all_gather_into_tensor(recv.view(torch.uint8), send.view(torch.uint8), group=self.group)
        recv = recv.view((self.world,) + pad_shape)
        parts = [recv[r, :, :sizes[r]] if axis == 0 else recv[r, :, :, :sizes[r]] for r in range(self.world)]
        full = torch.cat(parts, dim=1 if axis == 0 else 2).to(torch.int64)  # [K, B, T]
        return full.transpose(0, 1)

    def gather_codes_async(self, codes_local: torch.Tensor, global_shape: tp.Tuple[int, int, int]):
        """`gather_codes` on a side stream: returns (codes [B, K, T], event).  Nothing in the encode / decode of
        the next batch depends on the gathered codes, so the all-gather (and the int16 <-> int64 conversions around
        it) overlaps that work; a consumer calls `torch.cuda.current_stream().wait_event(event)` first."""
        if self.world == 1:
            ev = torch.cuda.Event()
            ev.record()
            return codes_local, ev
        dev = codes_local.device
        if getattr(self, "_gather_stream", None) is None:
            self._gather_stream = torch.cuda.Stream(dev)
        gs = self._gather_stream
        gs.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(gs):
            full = self.gather_codes(codes_local, global_shape)
            codes_local.record_stream(gs)
            ev = torch.cuda.Event()
            ev.record(gs)
        return full, ev

    # ---- whole-job calls ----------------------------------------------------------------------
    def encode(self, x_global_or_local: torch.Tensor, global_shape: tp.Optional[tp.Tuple[int, int, int]] = None,
               already_local: bool = False) -> torch.Tensor:
        """Encodes this rank's slice and returns the gathered codes of the whole batch.
        Pass the global tensor (it is sliced here) or, with already_local=True, this rank's slice plus
        the global [B, D, T] shape."""
        if already_local:
            assert global_shape is not None
            x_local, (B, _, T) = x_global_or_local, global_shape
        else:
            B, _, T = x_global_or_local.shape
            x_local = self.local_slice(x_global_or_local)
        codes_local = self.encode_local(x_local)
        return self.gather_codes(codes_local, (B, codes_local.shape[1], T))

    def decode(self, codes_global: torch.Tensor) -> torch.Tensor:
        """Decodes this rank's slice of gathered codes (local result; latents are not gathered)."""
        return self.decode_local(self.local_slice(codes_global))
