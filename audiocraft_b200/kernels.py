"""Torch-tensor front end of the C ABI (include/rvq_b200.h).

PyTorch is used here for device memory, the current stream and the device guard only; all
arithmetic happens in librvq_b200.so.  Every function requires CUDA tensors and raises otherwise:
there is deliberately no CPU or eager-PyTorch fallback.
"""
from __future__ import annotations

import ctypes
import typing as tp

import torch

from . import _lib

ALGO_AUTO, ALGO_EXACT, ALGO_TC = _lib.ALGO_AUTO, _lib.ALGO_EXACT, _lib.ALGO_TC
_ALGOS = {"auto": ALGO_AUTO, "exact": ALGO_EXACT, "tc": ALGO_TC}


def _require_cuda(t: torch.Tensor, what: str) -> None:
    if not t.is_cuda:
        raise RuntimeError(
            f"{what} must be a CUDA tensor: audiocraft_b200 runs the RVQ path on sm_100a kernels only "
            "(no CPU fallback)")


def _stream(device: torch.device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def launch_count(reset: bool = False) -> int:
    return int(_lib.load().rvq_b200_launch_count(1 if reset else 0))


def pack_codebooks(embeds: tp.Sequence[torch.Tensor]) -> torch.Tensor:
    """Builds the derived codebook pack (uint8 device tensor) from Ktot fp32 [C, D] codebooks."""
    lib = _lib.load()
    e0 = embeds[0]
    _require_cuda(e0, "codebooks")
    C, D = e0.shape
    Ktot = len(embeds)
    keep = []
    for e in embeds:
        if e.device != e0.device or e.dtype != torch.float32 or tuple(e.shape) != (C, D):
            raise ValueError("all codebooks must be fp32 [bins, dimension] tensors on one device")
        keep.append(e if e.is_contiguous() else e.contiguous())
    nbytes = lib.rvq_pack_bytes(Ktot, C, D)
    _lib.check(nbytes, "rvq_pack_bytes")
    with torch.cuda.device(e0.device):
        pack = torch.empty(nbytes, dtype=torch.uint8, device=e0.device)
        ptrs = (ctypes.c_void_p * Ktot)(*[e.data_ptr() for e in keep])
        rc = lib.rvq_pack_codebooks(ptrs, Ktot, C, D, pack.data_ptr(), nbytes, _stream(e0.device))
    _lib.check(rc, "rvq_pack_codebooks")
    return pack


def encode(x: torch.Tensor, pack: torch.Tensor, Ktot: int, C: int, K: int, want_quantized: bool = False,
           algo: str = "auto", wire: tp.Optional[torch.Tensor] = None
           ) -> tp.Tuple[torch.Tensor, tp.Optional[torch.Tensor]]:
    """x [B, D, T] fp32 (any strides) -> codes [K, B, T] int64 and, optionally, the summed
    quantized latent as a contiguous [B, T, D] tensor.  `wire`: an int16 [K, B, T] view (time axis contiguous, any
    k / b strides) that receives the same codes in the wire format of a frame-sharded job."""
    lib = _lib.load()
    _require_cuda(x, "x")
    if x.dtype != torch.float32:
        # the reference's matmul raises on mixed dtypes (core_vq.py:168)
        raise RuntimeError(f"expected a float32 input, got {x.dtype} (codebooks are float32)")
    if x.dim() != 3:
        raise ValueError(f"expected x of shape [B, D, T], got {tuple(x.shape)}")
    B, D, T = x.shape
    dev = x.device
    with torch.cuda.device(dev):
        codes = torch.empty((K, B, T), dtype=torch.int64, device=dev)
        quant = torch.empty((B, T, D), dtype=torch.float32, device=dev) if want_quantized else None
        if B * T > 0 and K > 0:
            a = _ALGOS[algo]
            ws_bytes = lib.rvq_encode_workspace_bytes(B, D, T, K, C, a)
            _lib.check(ws_bytes, "rvq_encode_workspace_bytes")
            ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev) if ws_bytes > 0 else None
            qp = quant.data_ptr() if quant is not None else None
            wp = ws.data_ptr() if ws is not None else None
            if wire is None:
                rc = lib.rvq_encode_f32(x.data_ptr(), _lib.strides3(x.stride()), B, D, T, pack.data_ptr(), Ktot, C, K,
                                        codes.data_ptr(), qp, wp, ws_bytes, a, _stream(dev))
            else:
                if wire.dtype != torch.int16 or tuple(wire.shape) != (K, B, T) or wire.device != dev or \
                        (T > 1 and wire.stride(2) != 1):
                    raise ValueError("wire must be an int16 [K, B, T] view on x's device with a contiguous time axis")
                rc = lib.rvq_encode_f32_wire(x.data_ptr(), _lib.strides3(x.stride()), B, D, T, pack.data_ptr(), Ktot, C,
                                             K, codes.data_ptr(), wire.data_ptr(),
                                             _lib._i64x2(int(wire.stride(0)), int(wire.stride(1))), qp, wp, ws_bytes, a,
                                             _stream(dev))
            _lib.check(rc, "rvq_encode_f32")
        elif quant is not None:
            quant.zero_()
    return codes, quant


def decode(codes: torch.Tensor, pack: torch.Tensor, Ktot: int, C: int, D: int, status: torch.Tensor,
           out: tp.Optional[torch.Tensor] = None, steps_late: tp.Optional[tp.Sequence[int]] = None,
           frames: tp.Optional[int] = None, allow_int16: bool = False) -> torch.Tensor:
    """codes [B, K, T] int64 / int32 (any strides) -> contiguous [B, T, D] fp32 (the reference's physical layout).
    int16 codes (the compact cache format) are accepted with allow_int16=True; without it they raise like the
    reference's F.embedding does.  `status` is the int64[2] error word of rvq_decode_f32.
    `steps_late` (with `frames`): `codes` is a delayed-pattern sequence [B, K, S] read in place; stage k's code of
    frame t is codes[b, k, t + steps_late[k]] and `frames` frames are decoded (rvq_decode_delayed_f32)."""
    lib = _lib.load()
    _require_cuda(codes, "codes")
    if codes.dtype not in ((torch.int64, torch.int32, torch.int16) if allow_int16 else (torch.int64, torch.int32)):
        raise RuntimeError(
            f"Expected tensor for argument #1 'indices' to have one of the following scalar types: Long, Int; "
            f"but got {codes.dtype} instead")
    if codes.dim() != 3:
        raise ValueError(f"expected codes of shape [B, K, T], got {tuple(codes.shape)}")
    B, K, T = codes.shape
    dev = codes.device
    if steps_late is not None:
        if frames is None or len(steps_late) != K or (K and frames + max(steps_late) > T) or frames < 0:
            raise ValueError("a delayed sequence needs K delays and a frame count with frames + max(delay) <= S")
        T = int(frames)
    with torch.cuda.device(dev):
        if out is None:
            out = torch.empty((B, T, D), dtype=torch.float32, device=dev)
        if B * T > 0:
            if steps_late is None:
                rc = lib.rvq_decode_f32(codes.data_ptr(), codes.element_size(), _lib.strides3(codes.stride()), B, K, T,
                                        pack.data_ptr(), Ktot, C, D, out.data_ptr(), _lib.strides3((T * D, 1, D)),
                                        status.data_ptr(), _stream(dev))
            else:
                late = (ctypes.c_int * max(K, 1))(*[int(v) for v in steps_late])
                rc = lib.rvq_decode_delayed_f32(codes.data_ptr(), codes.element_size(), _lib.strides3(codes.stride()), B, K,
                                                T, late, pack.data_ptr(), Ktot, C, D, out.data_ptr(),
                                                _lib.strides3((T * D, 1, D)), status.data_ptr(), _stream(dev))
            _lib.check(rc, "rvq_decode_f32")
    return out


# ---- code-stream plumbing (include/rvq_b200.h, csrc/rvq_codestream.cu) ------------------------------------------
def _require_codes(t: torch.Tensor, what: str, dtype=torch.int64) -> None:
    _require_cuda(t, what)
    if t.dtype != dtype:
        raise RuntimeError(f"{what} must be {dtype}, got {t.dtype}")


def widen_codes_i16(recv: torch.Tensor, W: int, K: int, B: int, T: int, axis: int,
                    out: tp.Optional[torch.Tensor] = None) -> torch.Tensor:
    """All-gathered int16 codes (W ranks x padded shard, see rvq_widen_codes_i16) -> int64 [K, B, T]."""
    lib = _lib.load()
    _require_codes(recv, "recv", torch.int16)
    dev = recv.device
    with torch.cuda.device(dev):
        if out is None:
            out = torch.empty((K, B, T), dtype=torch.int64, device=dev)
        elif out.dtype != torch.int64 or tuple(out.shape) != (K, B, T) or not out.is_contiguous():
            raise ValueError("out must be a contiguous int64 [K, B, T] tensor")
        rc = lib.rvq_widen_codes_i16(recv.data_ptr(), W, K, B, T, axis, out.data_ptr(), _stream(dev))
    _lib.check(rc, "rvq_widen_codes_i16")
    return out


_CODE_BYTES = {torch.int64: 8, torch.int32: 4, torch.int16: 2}


def convert_codes(codes: torch.Tensor, dtype: torch.dtype, out: tp.Optional[torch.Tensor] = None) -> torch.Tensor:
    """[B, K, T] codes (any strides; int64 / int32 / int16) -> contiguous [B, K, T] of `dtype`, or into `out`
    (same shape, any strides)."""
    lib = _lib.load()
    _require_cuda(codes, "codes")
    if codes.dtype not in _CODE_BYTES or dtype not in _CODE_BYTES or codes.dim() != 3:
        raise ValueError("codes must be a 3-d int64 / int32 / int16 tensor")
    B, K, T = codes.shape
    dev = codes.device
    with torch.cuda.device(dev):
        if out is None:
            out = torch.empty((B, K, T), dtype=dtype, device=dev)
        rc = lib.rvq_codes_convert(codes.data_ptr(), _CODE_BYTES[codes.dtype], _lib.strides3(codes.stride()), B, K, T,
                                   out.data_ptr(), _CODE_BYTES[out.dtype], _lib.strides3(out.stride()), _stream(dev))
    _lib.check(rc, "rvq_codes_convert")
    return out


def pack_codes(codes: torch.Tensor, bits: int) -> torch.Tensor:
    """Contiguous int64 codes (any shape) -> uint8 [ceil(n * bits / 8)]."""
    lib = _lib.load()
    _require_codes(codes, "codes")
    c = codes if codes.is_contiguous() else codes.contiguous()
    n = c.numel()
    nbytes = lib.rvq_codes_packed_bytes(n, bits)
    _lib.check(nbytes, "rvq_codes_packed_bytes")
    with torch.cuda.device(c.device):
        out = torch.empty(nbytes, dtype=torch.uint8, device=c.device)
        rc = lib.rvq_codes_pack_bits(c.data_ptr(), n, bits, out.data_ptr(), _stream(c.device))
    _lib.check(rc, "rvq_codes_pack_bits")
    return out


def unpack_codes(packed: torch.Tensor, shape: tp.Sequence[int], bits: int) -> torch.Tensor:
    lib = _lib.load()
    _require_codes(packed, "packed", torch.uint8)
    n = 1
    for d in shape:
        n *= int(d)
    if packed.numel() != lib.rvq_codes_packed_bytes(n, bits) or not packed.is_contiguous():
        raise ValueError("packed does not hold that many codes")
    with torch.cuda.device(packed.device):
        out = torch.empty(tuple(shape), dtype=torch.int64, device=packed.device)
        rc = lib.rvq_codes_unpack_bits(packed.data_ptr(), n, bits, out.data_ptr(), _stream(packed.device))
    _lib.check(rc, "rvq_codes_unpack_bits")
    return out


def stereo_interleave(c0: torch.Tensor, c1: torch.Tensor, per_timestep: bool) -> torch.Tensor:
    """Two [B, K, T] int64 code tensors (same strides) -> [B, 2K, T] or, per_timestep, [B, K, 2T]."""
    lib = _lib.load()
    _require_codes(c0, "codes of channel 0")
    _require_codes(c1, "codes of channel 1")
    if c0.shape != c1.shape or c0.dim() != 3:
        raise ValueError("the two channels must have the same [B, K, T] shape")
    if c1.stride() != c0.stride():
        c0, c1 = c0.contiguous(), c1.contiguous()
    B, K, T = c0.shape
    dev = c0.device
    with torch.cuda.device(dev):
        out = torch.empty((B, K, 2 * T) if per_timestep else (B, 2 * K, T), dtype=torch.int64, device=dev)
        rc = lib.rvq_stereo_interleave(c0.data_ptr(), c1.data_ptr(), _lib.strides3(c0.stride()), B, K, T,
                                       1 if per_timestep else 0, out.data_ptr(), _stream(dev))
    _lib.check(rc, "rvq_stereo_interleave")
    return out


def stereo_split(joint: torch.Tensor, per_timestep: bool) -> tp.Tuple[torch.Tensor, torch.Tensor]:
    lib = _lib.load()
    _require_codes(joint, "codes")
    if joint.dim() != 3 or joint.shape[2 if per_timestep else 1] % 2:
        raise ValueError("expected [B, 2K, T] codes (or [B, K, 2T] with per_timestep)")
    j = joint if joint.is_contiguous() else joint.contiguous()
    B = j.shape[0]
    K, T = (j.shape[1], j.shape[2] // 2) if per_timestep else (j.shape[1] // 2, j.shape[2])
    dev = j.device
    with torch.cuda.device(dev):
        o0 = torch.empty((B, K, T), dtype=torch.int64, device=dev)
        o1 = torch.empty((B, K, T), dtype=torch.int64, device=dev)
        rc = lib.rvq_stereo_split(j.data_ptr(), B, K, T, 1 if per_timestep else 0, o0.data_ptr(), o1.data_ptr(),
                                  _stream(dev))
    _lib.check(rc, "rvq_stereo_split")
    return o0, o1


def delay_pattern(seq: torch.Tensor, delays: tp.Sequence[int], offset: int, special: int, out_len: int,
                  revert: bool, valid_steps: tp.Optional[int] = None
                  ) -> tp.Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """build (revert=False): z [B, K, T] -> values [B, K, out_len]; revert: seq [B, K, S] -> values [B, K, out_len].
    Returns (values, indexes [K, out_len] int64, mask [K, out_len] bool) like the reference's Pattern methods."""
    lib = _lib.load()
    _require_codes(seq, "codes")
    if seq.dim() != 3 or len(delays) != seq.shape[1]:
        raise ValueError("expected [B, K, L] codes and K delays")
    B, K, L = seq.shape
    dev = seq.device
    d = (ctypes.c_int * K)(*[int(v) for v in delays])
    with torch.cuda.device(dev):
        values = torch.empty((B, K, out_len), dtype=torch.int64, device=dev)
        indexes = torch.empty((K, out_len), dtype=torch.int64, device=dev)
        mask = torch.empty((K, out_len), dtype=torch.uint8, device=dev)
        if revert:
            rc = lib.rvq_delay_revert(seq.data_ptr(), _lib.strides3(seq.stride()), B, K, L,
                                      L if valid_steps is None else int(valid_steps), out_len, d, int(offset), int(special),
                                      values.data_ptr(), indexes.data_ptr(), mask.data_ptr(), _stream(dev))
        else:
            rc = lib.rvq_delay_build(seq.data_ptr(), _lib.strides3(seq.stride()), B, K, L, out_len, d, int(offset),
                                     int(special), values.data_ptr(), indexes.data_ptr(), mask.data_ptr(), _stream(dev))
    _lib.check(rc, "rvq_delay_revert" if revert else "rvq_delay_build")
    return values, indexes, mask.bool()


# ---- peer-to-peer code exchange (include/rvq_b200.h, csrc/rvq_exchange.cu) ----------------------------------------
class ExchangeGeometry:
    """Byte layout of one rank's symmetric exchange buffer for codes [K, B, T] sharded over W ranks."""

    def __init__(self, W: int, K: int, B: int, T: int, nslots: int = 2):
        lib = _lib.load()
        self.W, self.K, self.B, self.T, self.nslots = W, K, B, T, nslots
        self.axis = 0 if B >= W else 2
        n = B if self.axis == 0 else T
        self.smax = n // W + (1 if n % W else 0)
        self.shard_shape = (K, self.smax, T) if self.axis == 0 else (K, B, self.smax)
        raw = 2 * K * self.smax * (T if self.axis == 0 else B)
        self.shard_bytes = (raw + 15) // 16 * 16
        self.slot_bytes = W * self.shard_bytes
        self.header = int(lib.rvq_exchange_header_bytes())
        self.total_bytes = self.header + nslots * self.slot_bytes

    def shard_view(self, buf: torch.Tensor, slot: int, rank: int) -> torch.Tensor:
        """int16 view [K, smax, T] / [K, B, smax] of rank `rank`'s shard of `slot` inside the uint8 buffer `buf`."""
        off = self.header + slot * self.slot_bytes + rank * self.shard_bytes
        n = self.shard_shape[0] * self.shard_shape[1] * self.shard_shape[2]
        return buf[off: off + 2 * n].view(torch.int16).view(self.shard_shape)


def exchange_push(geo: ExchangeGeometry, peer_ptrs: tp.Sequence[int], rank: int, seq: int, scratch: torch.Tensor,
                  status: torch.Tensor, stream: int, spin_wait: bool = False) -> None:
    lib = _lib.load()
    ptrs = (ctypes.c_void_p * geo.W)(*[int(p) for p in peer_ptrs])
    rc = lib.rvq_exchange_push(ptrs, geo.W, rank, seq % geo.nslots, geo.nslots, seq, geo.slot_bytes, geo.shard_bytes,
                               scratch.data_ptr(), status.data_ptr(), 1 if spin_wait else 0, stream)
    _lib.check(rc, "rvq_exchange_push")


def exchange_widen(geo: ExchangeGeometry, peer_ptrs: tp.Sequence[int], rank: int, seq: int, out: torch.Tensor,
                   scratch: torch.Tensor, status: torch.Tensor, stream: int, spin_wait: bool = False) -> None:
    lib = _lib.load()
    ptrs = (ctypes.c_void_p * geo.W)(*[int(p) for p in peer_ptrs])
    rc = lib.rvq_exchange_widen(ptrs, geo.W, rank, seq % geo.nslots, geo.nslots, seq, geo.slot_bytes, geo.shard_bytes,
                                geo.K, geo.B, geo.T, geo.axis, out.data_ptr(), scratch.data_ptr(), status.data_ptr(),
                                1 if spin_wait else 0, stream)
    _lib.check(rc, "rvq_exchange_widen")


# ---- training-time statistics (include/rvq_b200.h, csrc/rvq_train_stats.cu) ---------------------------------------
def ema_update(flat: torch.Tensor, codes: torch.Tensor, cluster_size: torch.Tensor, embed_avg: torch.Tensor,
               embed: torch.Tensor, decay: float, eps: float, want_stats: bool = False):
    """flat [N, D] fp32 (any strides), codes [N] int64: EMA step of cluster_size / embed_avg and the re-normalised embed,
    in place (core_vq.py:209-217).  Returns (counts [C], sums [C, D]) when asked."""
    lib = _lib.load()
    _require_cuda(flat, "x")
    for t, name in ((cluster_size, "cluster_size"), (embed_avg, "embed_avg"), (embed, "embed")):
        if not (t.is_cuda and t.is_contiguous() and t.dtype == torch.float32):
            raise RuntimeError(f"{name} must be a contiguous fp32 CUDA buffer")
    if flat.dtype != torch.float32 or flat.dim() != 2 or codes.dtype not in (torch.int64, torch.int32):
        raise RuntimeError("expected fp32 [N, D] inputs and int64 / int32 codes")
    N, D = flat.shape
    C = cluster_size.numel()
    c = codes.reshape(-1)
    if not c.is_contiguous():
        c = c.contiguous()
    dev = flat.device
    with torch.cuda.device(dev):
        counts = torch.empty(C, dtype=torch.float32, device=dev) if want_stats else None
        sums = torch.empty((C, D), dtype=torch.float32, device=dev) if want_stats else None
        rc = lib.rvq_ema_update_f32(flat.data_ptr(), _lib._i64x2(int(flat.stride(0)), int(flat.stride(1))), c.data_ptr(),
                                    c.element_size(), N, C, D, float(decay), float(eps), cluster_size.data_ptr(),
                                    embed_avg.data_ptr(), embed.data_ptr(), counts.data_ptr() if want_stats else None,
                                    sums.data_ptr() if want_stats else None, _stream(dev))
    _lib.check(rc, "rvq_ema_update_f32")
    return counts, sums
