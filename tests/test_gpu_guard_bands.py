"""Out-of-bounds writes: every buffer the kernels write (codes, quantized sum, decode output, encode workspace) is
carved out of a larger allocation whose margins hold a pattern; after the calls the margins must be untouched and
the results must equal those obtained with ordinary allocations.  Shapes are chosen so that the last tile, the
last 32-frame group and the last CTA are partial."""
import pytest
import torch

from audiocraft_b200 import _lib, kernels
from tests.golden import gen

pytestmark = pytest.mark.gpu

D = 128
GUARD = 1 << 16  # bytes on each side
PATTERN = 0xA5


def _guarded(nbytes, align=256):
    """(whole uint8 tensor, view of the payload): payload starts `GUARD` bytes in, aligned."""
    n = (nbytes + align - 1) // align * align
    whole = torch.full((GUARD + n + GUARD,), PATTERN, dtype=torch.uint8, device="cuda")
    assert whole.data_ptr() % align == 0
    return whole, whole[GUARD:GUARD + nbytes], nbytes


def _margins_intact(whole, nbytes):
    left = whole[:GUARD]
    right = whole[GUARD + nbytes:]
    return bool((left == PATTERN).all()) and bool((right == PATTERN).all())


@pytest.mark.parametrize("shape", [(4, 2048, 3, 701, "tc"), (2, 256, 5, 67, "tc"), (8, 1024, 1, 4099, "tc"),
                                   (4, 2048, 2, 333, "exact")])
def test_encode_and_decode_stay_inside_their_buffers(shape):
    K, C, B, T, algo = shape
    lib = _lib.load()
    embed = [torch.from_numpy(e).cuda() for e in gen.codebooks(600 + K, K, C, D)]
    pack = kernels.pack_codebooks(embed)
    x = torch.from_numpy(gen.normalish(700 + K, (B, D, T))).cuda()
    want_codes, want_quant = kernels.encode(x, pack, K, C, K, want_quantized=True, algo=algo)
    stream = torch.cuda.current_stream().cuda_stream
    algo_id = {"tc": _lib.ALGO_TC, "exact": _lib.ALGO_EXACT}[algo]

    ws_bytes = int(lib.rvq_encode_workspace_bytes(B, D, T, K, C, algo_id))
    ws_whole, ws, _ = _guarded(max(ws_bytes, 256))
    codes_whole, codes_raw, codes_bytes = _guarded(K * B * T * 8)
    quant_whole, quant_raw, quant_bytes = _guarded(B * T * D * 4)
    rc = lib.rvq_encode_f32(x.data_ptr(), _lib.strides3(x.stride()), B, D, T, pack.data_ptr(), K, C, K,
                            codes_raw.data_ptr(), quant_raw.data_ptr(), ws.data_ptr(), ws_bytes, algo_id, stream)
    assert rc == 0, _lib.last_error()
    torch.cuda.synchronize()
    assert _margins_intact(ws_whole, max(ws_bytes, 256)), "encode wrote outside its workspace"
    assert _margins_intact(codes_whole, codes_bytes), "encode wrote outside the codes"
    assert _margins_intact(quant_whole, quant_bytes), "encode wrote outside the quantized sum"
    codes = codes_raw.view(torch.int64).view(K, B, T)
    assert torch.equal(codes, want_codes)
    assert torch.equal(quant_raw.view(torch.float32).view(B, T, D).view(torch.int32), want_quant.view(torch.int32))

    # decode of those codes into a guarded [B, T, D] buffer
    out_whole, out_raw, out_bytes = _guarded(B * T * D * 4)
    status = torch.zeros(4, dtype=torch.int64, device="cuda")
    cs = _lib.strides3((T, B * T, 1))            # codes are [K, B, T]: strides of (b, k, t)
    os_ = _lib.strides3((T * D, 1, D))           # out is [B, T, D]: strides of (b, d, t)
    rc = lib.rvq_decode_f32(codes_raw.data_ptr(), 8, cs, B, K, T, pack.data_ptr(), K, C, D, out_raw.data_ptr(), os_,
                            status.data_ptr(), stream)
    assert rc == 0, _lib.last_error()
    torch.cuda.synchronize()
    assert _margins_intact(out_whole, out_bytes), "decode wrote outside its output"
    assert int(status[0]) == 0
    assert torch.equal(out_raw.view(torch.float32).view(B, T, D).view(torch.int32), want_quant.view(torch.int32))
