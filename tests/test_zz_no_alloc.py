"""rvq_encode_f32_static: the encode entry point for callers that cannot allocate device memory.

Runs last (file name) so that a failure here cannot hide the other tests behind `pytest -x`.  The new entry
point is called through the C ABI twice in a row (the second call finds the first call's leftovers in the
shared built-in scratch) and must return exactly what rvq_encode_f32 returns with a caller workspace:
same codes, same quantized sum, bit for bit -- the tolerance of every other parity test in this suite.
"""
import numpy as np
import pytest
import torch

from audiocraft_b200 import _lib, kernels
from tests.golden import gen

pytestmark = pytest.mark.gpu

K, C, D = 4, 2048, 128


def _pack():
    embed = [torch.from_numpy(e).cuda() for e in gen.codebooks(321, K, C, D)]
    return kernels.pack_codebooks(embed)


def _static_encode(lib, x, pack, want_quantized=True):
    """The caller owns inputs and outputs; the library brings its own scratch."""
    B, _, T = x.shape
    codes = torch.full((K, B, T), -1, dtype=torch.int64, device=x.device)
    quant = torch.full((B, T, D), float("nan"), device=x.device) if want_quantized else None
    rc = lib.rvq_encode_f32_static(x.data_ptr(), _lib.strides3(x.stride()), B, D, T, pack.data_ptr(), K, C, K,
                                   codes.data_ptr(), quant.data_ptr() if want_quantized else None,
                                   torch.cuda.current_stream().cuda_stream)
    return rc, codes, quant


def test_static_scratch_entry_point_matches_the_workspace_one_twice_in_a_row():
    lib = _lib.load()
    pack = _pack()
    # two different inputs of different sizes: the second call starts from a dirty scratch array
    shapes = [(3, 700), (1, 1500)]
    inputs = [torch.from_numpy(gen.normalish(400 + i, (b, D, t))).cuda() for i, (b, t) in enumerate(shapes)]
    for call, x in enumerate(inputs):
        want_codes, want_quant = kernels.encode(x, pack, K, C, K, want_quantized=True, algo="tc")
        rc, codes, quant = _static_encode(lib, x, pack)
        assert rc == 0, _lib.last_error()
        torch.cuda.synchronize()
        assert torch.equal(codes, want_codes), f"call {call}: codes differ"
        assert np.array_equal(quant.cpu().numpy().view(np.uint32), want_quant.cpu().numpy().view(np.uint32)), \
            f"call {call}: quantized sum differs"
    # codes only (no quantized output requested), immediately after
    rc, codes, _ = _static_encode(lib, inputs[0], pack, want_quantized=False)
    assert rc == 0, _lib.last_error()
    assert torch.equal(codes, kernels.encode(inputs[0], pack, K, C, K, algo="tc")[0])


def test_static_scratch_entry_point_takes_no_device_memory():
    """What the entry point is for: once the first (warm-up) call on the device has run, a call neither asks the
    CUDA allocator nor torch's for anything.  Free device memory, as the driver reports it, and torch's own
    counters must read the same before and after."""
    lib = _lib.load()
    pack = _pack()
    x = torch.from_numpy(gen.normalish(55, (2, D, 500))).cuda()
    B, _, T = x.shape
    codes = torch.empty((K, B, T), dtype=torch.int64, device="cuda")
    quant = torch.empty((B, T, D), device="cuda")
    args = (x.data_ptr(), _lib.strides3(x.stride()), B, D, T, pack.data_ptr(), K, C, K, codes.data_ptr(), quant.data_ptr(),
            torch.cuda.current_stream().cuda_stream)
    assert lib.rvq_encode_f32_static(*args) == 0, _lib.last_error()   # warm-up: module load + symbol look-up
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    alloc0, reserved0 = torch.cuda.memory_allocated(), torch.cuda.memory_reserved()
    for _ in range(3):
        assert lib.rvq_encode_f32_static(*args) == 0, _lib.last_error()
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert (torch.cuda.memory_allocated(), torch.cuda.memory_reserved()) == (alloc0, reserved0)
    assert free1 == free0, f"device memory changed by {free0 - free1} bytes across calls that must not allocate"


def test_static_scratch_entry_point_refuses_what_does_not_fit():
    lib = _lib.load()
    pack = _pack()
    x = torch.zeros((1, D, 4097), device="cuda")  # one frame more than RVQ_STATIC_MAX_FRAMES
    rc, _, _ = _static_encode(lib, x, pack, want_quantized=False)
    assert rc == _lib.RVQ_EINVAL
    assert "4096" in _lib.last_error() and "frames" in _lib.last_error()
    # shapes that never use scratch are pointed at the exact scan instead
    rc = lib.rvq_encode_f32_static(x.data_ptr(), _lib.strides3((1, 1, 1)), 1, 32, 8, pack.data_ptr(), K, 64, K,
                                   x.data_ptr(), None, None)
    assert rc == _lib.RVQ_EUNSUPPORTED
    # the limit itself is accepted
    x = torch.from_numpy(gen.normalish(77, (2, D, 2048))).cuda()
    rc, codes, _ = _static_encode(lib, x, pack, want_quantized=False)
    assert rc == 0, _lib.last_error()
    assert torch.equal(codes, kernels.encode(x, pack, K, C, K, algo="tc")[0])
