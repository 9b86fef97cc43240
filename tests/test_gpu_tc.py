"""GPU tests of the tensor-core (tcgen05) encode path: the approximate scores must stay inside the
error bound that certifies the exact re-scoring, and the emitted codes must equal the oracle's."""
import numpy as np
import pytest
import torch

import oracle
from audiocraft_b200 import _lib, kernels
from tests.golden import gen
from tests.test_gpu_parity import make_module, same_bits

pytestmark = pytest.mark.gpu


def tc_scores(m, x, stage):
    """Runs rvq_debug_tc_scores; returns codes [K,B,T], acc [128,C], info [128,8]."""
    lib = _lib.load()
    vq = m.vq
    Ktot, C, D = vq._shape()
    pack = vq._pack.get(vq._embeds())
    B, _, T = x.shape
    K = m.n_q
    codes = torch.empty((K, B, T), dtype=torch.int64, device=x.device)
    acc = torch.full((128, C), float("nan"), device=x.device)
    info = torch.zeros((128 * 8 + 192,), device=x.device)   # 128 x 8 fp32 + 96 int64 cycle counters
    ws_bytes = lib.rvq_encode_workspace_bytes(B, D, T, K, C, _lib.ALGO_TC)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=x.device)
    rc = lib.rvq_debug_tc_scores(x.data_ptr(), _lib.strides3(x.stride()), B, D, T, pack.data_ptr(), Ktot, C, K,
                                 stage, codes.data_ptr(), acc.data_ptr(), info.data_ptr(), ws.data_ptr(), ws_bytes,
                                 torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "rvq_debug_tc_scores")
    torch.cuda.synchronize()
    return codes.cpu().numpy(), acc.cpu().numpy(), info[:1024].view(128, 8).cpu().numpy()


@pytest.mark.parametrize("stage", [0, 3])
def test_tc_scores_are_within_the_certified_bound(stage):
    embed = gen.codebooks(7, 4, 2048, 128)
    x_np = gen.normalish(77, (2, 128, 300))
    m = make_module(embed, 128, 2048)
    codes, acc, info = tc_scores(m, torch.from_numpy(x_np).cuda(), stage)
    want = oracle.encode(x_np, embed).transpose(1, 0, 2)          # [K, B, T]
    # residual of the first 128 frames (batch item 0) entering `stage`, from the oracle's codes
    r = x_np[0, :, :128].T.astype(np.float32).copy()               # [128, D]
    for k in range(stage):
        r = r - embed[k][want[k, 0, :128]]
    e = embed[stage].astype(np.float64)
    exact = r.astype(np.float64) @ e.T - 0.5 * (e ** 2).sum(1)[None, :]   # r.c - cc/2
    scale = info[:, 6].astype(np.float64)[:, None]
    err = np.abs(acc.astype(np.float64) - scale * exact)
    delta = 0.5 * info[:, 2].astype(np.float64)[:, None]
    assert np.isfinite(acc).all()
    assert (info[:, 5] == 0).all(), "no frame of this input should need the fallback"
    ratio = (err / delta).max()
    print(f"stage {stage}: max |acc - exact| / delta = {ratio:.4f}; mean remembered groups {info[:, 4].mean():.2f}")
    assert ratio < 1.0
    assert ratio > 1e-4, "suspiciously exact: is the fp16 path really used?"
    assert np.array_equal(codes, want)


@pytest.mark.parametrize("shape", [(4, 2048, 3, 700), (32, 1024, 2, 301), (8, 2048, 1, 129), (2, 256, 5, 64),
                                   (3, 2048, 2, 333), (1, 512, 7, 100)])  # odd stage counts: the quantized sum walks stages in pairs
def test_tc_codes_equal_oracle(shape):
    K, C, B, T = shape
    embed = gen.codebooks(100 + K, K, C, 128)
    x_np = gen.normalish(200 + K, (B, 128, T))
    m = make_module(embed, 128, C)
    m.vq.encode_algo = "tc"
    with torch.no_grad():
        res = m(torch.from_numpy(x_np).cuda(), 50)
    o_q, o_codes = oracle.forward(x_np, embed)
    got = res.codes.cpu().numpy()
    assert np.array_equal(got, o_codes), f"{(got != o_codes).sum()} of {got.size} codes differ"
    assert same_bits(np.ascontiguousarray(res.x.cpu().numpy()), np.ascontiguousarray(o_q))


def test_tc_frame_major_input_and_many_tiles():
    """[B, T, D]-physical input (the style conditioner's layout) and more tiles than CTA slots."""
    embed = gen.codebooks(9, 2, 512, 128)
    x_btd = gen.normalish(91, (3, 14000, 128))
    x_np = x_btd.transpose(0, 2, 1)
    m = make_module(embed, 128, 512)
    m.vq.encode_algo = "tc"
    with torch.no_grad():
        codes = m.encode(torch.from_numpy(x_btd).cuda().transpose(1, 2))
    assert np.array_equal(codes.cpu().numpy(), oracle.encode(x_np, embed))


@pytest.mark.parametrize("name", ["nonfinite_frames", "tiny_scale_inputs", "large_scale_inputs", "zero_input",
                                  "dup_rows", "exact_rows_and_midpoints", "trained_like"])
def test_tc_fallback_and_tie_cases(name):
    case = gen.build_case(name)
    cfg = case["cfg"]
    m = make_module(case["embed"], cfg["D"], cfg["C"], case["n_q"])
    m.vq.encode_algo = "tc"
    with torch.no_grad():
        codes = m.encode(torch.from_numpy(case["x"]).cuda())
    assert np.array_equal(codes.cpu().numpy(), oracle.encode(case["x"], case["embed"], case["n_q"]))


def test_tc_rejects_uncovered_shapes():
    m = make_module(gen.codebooks(3, 2, 64, 32), 32, 64)
    m.vq.encode_algo = "tc"
    with pytest.raises(NotImplementedError):
        m.encode(torch.zeros(1, 32, 8, device="cuda"))
    m.vq.encode_algo = "auto"   # falls back to the exact scan
    assert m.encode(torch.zeros(1, 32, 8, device="cuda")).shape == (1, 2, 8)
