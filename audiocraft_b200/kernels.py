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
           algo: str = "auto") -> tp.Tuple[torch.Tensor, tp.Optional[torch.Tensor]]:
    """x [B, D, T] fp32 (any strides) -> codes [K, B, T] int64 and, optionally, the summed
    quantized latent as a contiguous [B, T, D] tensor."""
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
            rc = lib.rvq_encode_f32(x.data_ptr(), _lib.strides3(x.stride()), B, D, T, pack.data_ptr(), Ktot, C, K,
                                    codes.data_ptr(), quant.data_ptr() if quant is not None else None,
                                    ws.data_ptr() if ws is not None else None, ws_bytes, a, _stream(dev))
            _lib.check(rc, "rvq_encode_f32")
        elif quant is not None:
            quant.zero_()
    return codes, quant


def decode(codes: torch.Tensor, pack: torch.Tensor, Ktot: int, C: int, D: int, status: torch.Tensor,
           out: tp.Optional[torch.Tensor] = None) -> torch.Tensor:
    """codes [B, K, T] int64/int32 (any strides) -> contiguous [B, T, D] fp32 (the reference's
    physical layout).  `status` is the int64[2] error word of rvq_decode_f32."""
    lib = _lib.load()
    _require_cuda(codes, "codes")
    if codes.dtype not in (torch.int64, torch.int32):
        raise RuntimeError(
            f"Expected tensor for argument #1 'indices' to have one of the following scalar types: Long, Int; "
            f"but got {codes.dtype} instead")
    if codes.dim() != 3:
        raise ValueError(f"expected codes of shape [B, K, T], got {tuple(codes.shape)}")
    B, K, T = codes.shape
    dev = codes.device
    with torch.cuda.device(dev):
        if out is None:
            out = torch.empty((B, T, D), dtype=torch.float32, device=dev)
        if B * T > 0:
            rc = lib.rvq_decode_f32(codes.data_ptr(), codes.element_size(), _lib.strides3(codes.stride()), B, K, T,
                                    pack.data_ptr(), Ktot, C, D, out.data_ptr(), _lib.strides3((T * D, 1, D)),
                                    status.data_ptr(), _stream(dev))
            _lib.check(rc, "rvq_decode_f32")
    return out
