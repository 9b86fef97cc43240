"""ctypes binding of librvq_b200.so (include/rvq_b200.h).

There is no CPU fallback and no alternative backend: if the shared library is missing or does
not export the whole ABI, importing the kernels fails loudly.
"""
from __future__ import annotations

import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "lib", "librvq_b200.so")

RVQ_OK = 0
RVQ_EINVAL = -1
RVQ_EUNSUPPORTED = -2
RVQ_ECUDA = -3
RVQ_EPACK = -4

ALGO_AUTO, ALGO_EXACT, ALGO_TC = 0, 1, 2
ABI_VERSION = 2

_i32, _i64, _vp = ctypes.c_int, ctypes.c_int64, ctypes.c_void_p
_i64x3 = ctypes.c_int64 * 3
_i64x2 = ctypes.c_int64 * 2

# name -> (restype, argtypes); must list every symbol include/rvq_b200.h declares
SIGNATURES = {
    "rvq_b200_abi_version": (_i32, []),
    "rvq_b200_last_error": (ctypes.c_char_p, []),
    "rvq_b200_launch_count": (_i64, [_i32]),
    "rvq_pack_bytes": (_i64, [_i32, _i32, _i32]),
    "rvq_pack_codebooks": (_i32, [ctypes.POINTER(_vp), _i32, _i32, _i32, _vp, _i64, _vp]),
    "rvq_encode_workspace_bytes": (_i64, [_i32, _i32, _i32, _i32, _i32, _i32]),
    "rvq_encode_f32": (_i32, [_vp, _i64x3, _i32, _i32, _i32, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _i64,
                              _i32, _vp]),
    "rvq_encode_f32_wire": (_i32, [_vp, _i64x3, _i32, _i32, _i32, _vp, _i32, _i32, _i32, _vp, _vp, _i64x2, _vp, _vp, _i64,
                                   _i32, _vp]),
    "rvq_widen_codes_i16": (_i32, [_vp, _i32, _i32, _i32, _i32, _i32, _vp, _vp]),
    "rvq_exchange_header_bytes": (_i64, []),
    "rvq_debug_gate": (_i32, [_vp, _i64, _vp]),
    "rvq_b200_set_reserved_sms": (_i32, [_i32]),
    "rvq_exchange_push": (_i32, [ctypes.POINTER(_vp), _i32, _i32, _i32, _i32, ctypes.c_uint32, _i64, _i64, _vp, _vp, _i32,
                                 _vp]),
    "rvq_exchange_widen": (_i32, [ctypes.POINTER(_vp), _i32, _i32, _i32, _i32, ctypes.c_uint32, _i64, _i64, _i32, _i32, _i32,
                                  _i32, _vp, _vp, _vp, _i32, _vp]),
    "rvq_codes_convert": (_i32, [_vp, _i32, _i64x3, _i32, _i32, _i32, _vp, _i32, _i64x3, _vp]),
    "rvq_codes_packed_bytes": (_i64, [_i64, _i32]),
    "rvq_codes_pack_bits": (_i32, [_vp, _i64, _i32, _vp, _vp]),
    "rvq_codes_unpack_bits": (_i32, [_vp, _i64, _i32, _vp, _vp]),
    "rvq_stereo_interleave": (_i32, [_vp, _vp, _i64x3, _i32, _i32, _i32, _i32, _vp, _vp]),
    "rvq_stereo_split": (_i32, [_vp, _i32, _i32, _i32, _i32, _vp, _vp, _vp]),
    "rvq_delay_build": (_i32, [_vp, _i64x3, _i32, _i32, _i32, _i32, ctypes.POINTER(_i32), _i32, _i64, _vp, _vp, _vp, _vp]),
    "rvq_delay_revert": (_i32, [_vp, _i64x3, _i32, _i32, _i32, _i32, _i32, ctypes.POINTER(_i32), _i32, _i64, _vp, _vp,
                                _vp, _vp]),
    "rvq_encode_f32_static": (_i32, [_vp, _i64x3, _i32, _i32, _i32, _vp, _i32, _i32, _i32, _vp, _vp, _vp]),
    "rvq_debug_tc_scores": (_i32, [_vp, _i64x3, _i32, _i32, _i32, _vp, _i32, _i32, _i32, _i32, _vp, _vp, _vp, _vp,
                                   _i64, _vp]),
    "rvq_decode_f32": (_i32, [_vp, _i32, _i64x3, _i32, _i32, _i32, _vp, _i32, _i32, _i32, _vp, _i64x3, _vp,
                              _vp]),
    "rvq_ema_update_f32": (_i32, [_vp, _i64x2, _vp, _i32, _i64, _i32, _i32, ctypes.c_float, ctypes.c_float, _vp, _vp, _vp, _vp,
                                  _vp, _vp]),
    "rvq_decode_delayed_f32": (_i32, [_vp, _i32, _i64x3, _i32, _i32, _i32, ctypes.POINTER(_i32), _vp, _i32, _i32, _i32, _vp,
                                      _i64x3, _vp, _vp]),
}

_lib = None


class RvqLibraryError(RuntimeError):
    pass


def load() -> ctypes.CDLL:
    """Loads the library once; raises RvqLibraryError when it is absent or incomplete."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RvqLibraryError(
            f"{LIB_PATH} is missing: build it with `python -m audiocraft_b200.build` "
            "(there is no CPU or PyTorch fallback for the RVQ kernels)")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (restype, argtypes) in SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as e:
            raise RvqLibraryError(f"{LIB_PATH} does not export {name}") from e
        fn.restype = restype
        fn.argtypes = argtypes
    ver = lib.rvq_b200_abi_version()
    if ver != ABI_VERSION:
        raise RvqLibraryError(f"ABI version mismatch: library {ver}, binding {ABI_VERSION}")
    _lib = lib
    return lib


def last_error() -> str:
    return load().rvq_b200_last_error().decode("utf-8", "replace")


def check(rc: int, what: str) -> None:
    """Maps a negative return code to the exception the reference would raise."""
    if rc >= 0:
        return
    msg = f"{what}: {last_error()}"
    if rc == RVQ_EUNSUPPORTED:
        raise NotImplementedError(msg)
    if rc == RVQ_EINVAL:
        raise ValueError(msg)
    raise RuntimeError(msg)


def strides3(values) -> "ctypes.Array":
    return _i64x3(*[int(v) for v in values])
