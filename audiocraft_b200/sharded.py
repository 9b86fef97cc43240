"""Frame-sharded RVQ over the GPUs of one box (one process per GPU, torch.distributed / NCCL).

The reference's inference path is single-device; this is the multi-GPU design of
BASELINE.json config 5.  RVQ is frame independent -- the K-stage chain is sequential only inside a
frame -- so any split of the B*T frames is valid: no halo, no exchange during compute.

  * codebooks: replicated, one broadcast of the module's buffers from rank 0 at construction
  * compute:   every rank encodes / decodes its contiguous slice (whole batch items when B >= world,
               else a slice of the time axis)
  * exchange:  one all-gather of the codes per call, as int16 on the wire when bins <= 32767
               (4x fewer bytes than int64).  On CUDA the encode kernel writes the int16 copy straight
               into a preallocated send buffer and one kernel of librvq_b200.so turns the gathered
               buffer into the [K, B, T] int64 block: two launches per call beside the collective.  Quantized latents are not gathered:
               re-decoding gathered codes locally is cheaper than moving fp32 latents over NVLink.
"""
from __future__ import annotations

import os
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

    def __init__(self, quantizer, group: tp.Optional[dist.ProcessGroup] = None, wire_int16: tp.Optional[bool] = None,
                 exchange: str = "auto"):
        self.q = quantizer
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        bins = getattr(quantizer, "bins", 1 << 30)
        self.wire_int16 = (bins <= 32767) if wire_int16 is None else wire_int16
        self._wire_sets: tp.Dict[tuple, dict] = {}
        self._wired = None
        # "p2p": peer stores over NVLink into symmetric memory + flags (csrc/rvq_exchange.cu); "nccl": all_gather;
        # "auto": p2p when torch's symmetric memory can be set up for the group, else nccl
        assert exchange in ("auto", "p2p", "nccl")
        self.exchange = exchange
        self._p2p: tp.Dict[tuple, tp.Any] = {}
        if self.world > 1 and torch.cuda.is_available():
            # the persistent encode kernel leaves a few SMs to the exchange kernels (config 2's 750 tiles take three
            # rounds on 140 CTAs as on 148: nothing is lost); more ranks move more bytes per step
            from . import _lib
            n = int(os.environ.get("RVQ_RESERVED_SMS", "4" if self.world <= 2 else "8"))
            _lib.check(_lib.load().rvq_b200_set_reserved_sms(n), "rvq_b200_set_reserved_sms")
        if self.world > 1 and hasattr(quantizer, "buffers"):
            for buf in quantizer.buffers():
                dist.broadcast(buf, src=0, group=group)
            vq = getattr(quantizer, "vq", None)
            if vq is not None and hasattr(vq, "invalidate_cache"):
                vq.invalidate_cache()  # the derived codebook pack must be rebuilt from rank 0's buffers

    # ---- local pieces -----------------------------------------------------------------------
    def local_slice(self, x: torch.Tensor) -> torch.Tensor:
        """This rank's part of a global [B, D, T] (or codes [B, K, T]) tensor."""
        B, _, T = x.shape
        axis = shard_axis(B, T, self.world)
        lo, hi = shard_bounds(x.shape[axis], self.world, self.rank)
        return x[lo:hi] if axis == 0 else x[:, :, lo:hi]

    def encode_local(self, x_local: torch.Tensor, global_shape: tp.Optional[tp.Tuple[int, int, int]] = None) -> torch.Tensor:
        """Encodes this rank's slice.  With `global_shape` ([B, D or K, T] of the whole batch) on a CUDA device the
        encode kernel also writes the codes as int16 straight into the send buffer of the coming all-gather
        (`rvq_encode_f32_wire`), so `gather_codes` of the returned tensor needs no narrowing pass."""
        vq = getattr(self.q, "vq", None)
        if global_shape is None or self.world == 1 or not x_local.is_cuda or not self.wire_int16 or vq is None:
            return self.q.encode(x_local)
        B, _, T = global_shape
        K = self.q.n_q
        bufs = self._wire_buffers(K, B, T, x_local.device)
        cur = torch.cuda.current_stream(x_local.device)
        if bufs["free"] is not None:
            cur.wait_event(bufs["free"])          # the all-gather that read this send buffer last
        b_loc, t_loc = x_local.shape[0], x_local.shape[2]
        wire = bufs["send"][:, :b_loc, :t_loc]
        codes = vq.encode(x_local, n_q=K, wire=wire).transpose(0, 1)
        self._wired = (codes, bufs)
        return codes

    def decode_local(self, codes_local: torch.Tensor) -> torch.Tensor:
        return self.q.decode(codes_local)

    # ---- exchange ---------------------------------------------------------------------------
    def _wire_buffers(self, K: int, B: int, T: int, device: torch.device) -> dict:
        """Send / receive buffers of the int16 all-gather, allocated once per shape, two sets used in turn (the
        all-gather of one batch runs beside the encode of the next)."""
        key = (K, B, T, device)
        p2p = self._p2p_state(K, B, T, device)
        if p2p is not None:
            p2p["seq"] += 1
            seq = p2p["seq"]
            geo = p2p["geo"]
            # "free": my push and widen of the step that used this slot last (they read my shard of the slot)
            return {"send": geo.shard_view(p2p["buf"], seq % geo.nslots, self.rank), "free": p2p["free"][seq % geo.nslots],
                    "axis": geo.axis, "p2p": p2p, "seq": seq}
        sets = self._wire_sets.get(key)
        if sets is None:
            axis = shard_axis(B, T, self.world)
            smax = max(shard_sizes(B if axis == 0 else T, self.world))
            shape = (K, smax, T) if axis == 0 else (K, B, smax)
            sets = {"next": 0, "sets": [
                {"send": torch.zeros(shape, dtype=torch.int16, device=device),
                 "recv": torch.empty((self.world,) + shape, dtype=torch.int16, device=device),
                 "free": None, "axis": axis} for _ in range(2)]}
            self._wire_sets[key] = sets
        cur = sets["sets"][sets["next"]]
        sets["next"] ^= 1
        return cur

    def _p2p_state(self, K: int, B: int, T: int, device: torch.device):
        """Symmetric exchange buffer for this shape (allocated and rendezvoused once, collectively), or None."""
        if self.exchange == "nccl" or self.world == 1 or device.type != "cuda" or not self.wire_int16:
            return None
        key = (K, B, T, device)
        if key in self._p2p:
            return self._p2p[key]
        from . import kernels
        state = None
        try:
            import torch.distributed._symmetric_memory as symm_mem
            # slots = how many steps the exchange may lag behind the encodes before an encode has to wait for it: as many
            # as fit 256 MB (and the flag area), at least two
            probe = kernels.ExchangeGeometry(self.world, K, B, T, nslots=2)
            nslots = int(max(2, min(8, 128 // self.world, (256 << 20) // max(probe.slot_bytes, 1))))
            geo = kernels.ExchangeGeometry(self.world, K, B, T, nslots=nslots)
            buf = symm_mem.empty(geo.total_bytes, dtype=torch.uint8, device=device)
            hdl = symm_mem.rendezvous(buf, self.group if self.group is not None else dist.group.WORLD)
            buf.zero_()
            torch.cuda.synchronize(device)
            dist.barrier(group=self.group)     # nobody pushes into a buffer that is still being cleared
            state = {"geo": geo, "buf": buf, "hdl": hdl, "ptrs": [int(p) for p in hdl.buffer_ptrs], "seq": 0, "free": [None] * nslots,
                     "scratch": torch.zeros(256, dtype=torch.uint8, device=device),
                     "status": torch.zeros(2, dtype=torch.int64, device=device)}
        except Exception as e:  # symmetric memory not available for this group / driver
            if self.exchange == "p2p":
                raise
            import warnings
            warnings.warn(f"symmetric memory unavailable ({e}); the codes are exchanged with NCCL all_gather")
        # every rank must take the same path: agree
        flag = torch.tensor([1 if state is not None else 0], device=device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.group)
        if int(flag.item()) == 0:
            state = None
        self._p2p[key] = state
        return state

    def check_exchange(self) -> None:
        """Raises if a wait of the peer-to-peer exchange timed out since the last check (a rank did not show up)."""
        for st in self._p2p.values():
            if st is not None:
                what, who = st["status"].tolist()
                if what:
                    st["status"].zero_()
                    raise RuntimeError(f"code exchange: {'push' if what == 1 else 'widen'} timed out waiting for rank {who}")

    def gather_codes(self, codes_local: torch.Tensor, global_shape: tp.Tuple[int, int, int]) -> torch.Tensor:
        """All-gathers per-rank codes [B_r, K, T_r] into the global [B, K, T] (int64) on every rank."""
        B, K, T = global_shape
        if self.world == 1:
            return codes_local
        if codes_local.is_cuda and self.wire_int16:
            return self._gather_codes_cuda(codes_local, B, K, T)
        # host-logic path (CPU tensors under gloo in the tests; int64 on the wire for huge codebooks)
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

    def _gather_codes_cuda(self, codes_local: torch.Tensor, B: int, K: int, T: int,
                           out: tp.Optional[torch.Tensor] = None) -> torch.Tensor:
        """One all-gather of the int16 send buffer + one widening kernel (`rvq_widen_codes_i16`)."""
        from . import kernels
        wired = getattr(self, "_wired", None)
        if wired is not None and wired[0] is codes_local:
            bufs = wired[1]                       # the encode kernel already filled the send buffer
        else:
            bufs = self._wire_buffers(K, B, T, codes_local.device)
            cur = torch.cuda.current_stream(codes_local.device)
            if bufs["free"] is not None:
                cur.wait_event(bufs["free"])
            b_loc, t_loc = codes_local.shape[0], codes_local.shape[2]
            kernels.convert_codes(codes_local, torch.int16, out=bufs["send"][:, :b_loc, :t_loc].transpose(0, 1))
        self._wired = None
        if "p2p" in bufs:
            st = bufs["p2p"]
            dev = codes_local.device
            if out is None:
                out = torch.empty((K, B, T), dtype=torch.int64, device=dev)
            stream = torch.cuda.current_stream(dev).cuda_stream
            # in-kernel waits by default: the exchange kernels run on the SMs the encode kernel leaves free and notice a
            # flag within microseconds; RVQ_EXCHANGE_STREAM_WAIT=1 selects stream memory operations instead
            spin = st.get("spin", os.environ.get("RVQ_EXCHANGE_STREAM_WAIT") != "1")
            trace = getattr(self, "trace", None)
            if trace is not None:
                e0 = torch.cuda.Event(enable_timing=True); e0.record(torch.cuda.current_stream(dev))
            try:
                kernels.exchange_push(st["geo"], st["ptrs"], self.rank, bufs["seq"], st["scratch"], st["status"], stream,
                                      spin_wait=spin)
            except RuntimeError:
                if spin:
                    raise
                # stream memory operations refused (driver / memory type): wait inside the kernels instead
                spin = st["spin"] = True
                kernels.exchange_push(st["geo"], st["ptrs"], self.rank, bufs["seq"], st["scratch"], st["status"], stream,
                                      spin_wait=True)
            if trace is not None:
                e1 = torch.cuda.Event(enable_timing=True); e1.record(torch.cuda.current_stream(dev))
            kernels.exchange_widen(st["geo"], st["ptrs"], self.rank, bufs["seq"], out, st["scratch"], st["status"], stream,
                                   spin_wait=spin)
            if trace is not None:
                e2 = torch.cuda.Event(enable_timing=True); e2.record(torch.cuda.current_stream(dev))
                trace.append((e0, e1, e2))
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(dev))
            st["free"][bufs["seq"] % st["geo"].nslots] = ev
            return out.transpose(0, 1)
        dist.all_gather_into_tensor(bufs["recv"].view(torch.uint8).view(-1), bufs["send"].view(torch.uint8).view(-1),
                                    group=self.group)
        full = kernels.widen_codes_i16(bufs["recv"], self.world, K, B, T, bufs["axis"], out=out)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(codes_local.device))
        bufs["free"] = ev
        return full.transpose(0, 1)

    def gather_codes_async(self, codes_local: torch.Tensor, global_shape: tp.Tuple[int, int, int]):
        """`gather_codes` on a side stream: returns (codes [B, K, T], event).  Nothing in the encode / decode of
        the next batch depends on the gathered codes, so the all-gather and the widening kernel overlap that work;
        a consumer calls `torch.cuda.current_stream().wait_event(event)` first."""
        if self.world == 1:
            ev = torch.cuda.Event()
            ev.record()
            return codes_local, ev
        dev = codes_local.device
        if getattr(self, "_gather_stream", None) is None:
            # high priority: when SMs free up between two persistent encode kernels, the exchange's few small blocks
            # go first instead of waiting a whole encode for room
            self._gather_stream = torch.cuda.Stream(dev, priority=-1)
        gs = self._gather_stream
        cur = torch.cuda.current_stream(dev)
        B, K, T = global_shape
        fast = codes_local.is_cuda and self.wire_int16
        wired = fast and self._wired is not None and self._wired[0] is codes_local
        # the result is allocated on the caller's stream (its pool is the warm one) and filled on the gather stream
        out = torch.empty((K, B, T), dtype=torch.int64, device=dev) if fast else None
        gs.wait_stream(cur)
        with torch.cuda.stream(gs):
            if fast:
                full = self._gather_codes_cuda(codes_local, B, K, T, out=out)
                out.record_stream(gs)
            else:
                full = self.gather_codes(codes_local, global_shape)
                full.record_stream(cur)
            if not wired:
                codes_local.record_stream(gs)  # read on the gather stream (the wired path reads the send buffer only)
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
        K = getattr(self.q, "n_q", None)
        codes_local = self.encode_local(x_local, (B, K, T) if K is not None else None)
        return self.gather_codes(codes_local, (B, codes_local.shape[1], T))

    def decode(self, codes_global: torch.Tensor) -> torch.Tensor:
        """Decodes this rank's slice of gathered codes (local result; latents are not gathered)."""
        return self.decode_local(self.local_slice(codes_global))


def average_buffers(module, group: tp.Optional[dist.ProcessGroup] = None) -> None:
    """Data-parallel training: the quantizer's buffers (`inited`, `cluster_size`, `embed`, `embed_avg` of every
    layer) become their mean over the workers, as `flashy.distrib.average_tensors(self.rvq.buffers())` does after a
    training step of the style conditioner (modules/conditioners.py:963).  One all-reduce over one flat fp32 buffer
    per call instead of one collective per tensor (4 per layer); the derived codebook pack is invalidated."""
    if not (dist.is_available() and dist.is_initialized()):
        return
    world = dist.get_world_size(group)
    if world == 1:
        return
    bufs = [b for b in module.buffers() if b.is_floating_point()]
    if not bufs:
        return
    flat = torch.cat([b.detach().reshape(-1).to(torch.float32) for b in bufs])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    flat.div_(world)
    off = 0
    with torch.no_grad():
        for b in bufs:
            n = b.numel()
            b.copy_(flat[off:off + n].view_as(b))
            off += n
    vq = getattr(module, "vq", None)
    if vq is not None and hasattr(vq, "invalidate_cache"):
        vq.invalidate_cache()
