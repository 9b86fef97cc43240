"""GPU parity tests: the CUDA path, called through the drop-in module and the C ABI, against the CPU
oracle (bit exact) and against the golden vectors of the unmodified reference.

Bar: codes bit-exact vs the oracle on every case; vs the reference's golden codes exact except fp32
near-ties adjudicated in fp64 (tests/parity.py); dequantized output bit-exact (same gather, same fp32
summation order -- no tolerance needed).
"""
import os

import numpy as np
import pytest
import torch

import oracle
from audiocraft_b200 import kernels
from audiocraft_b200.quantization import QuantizedResult, ResidualVectorQuantizer
from tests.golden import gen
from tests.parity import adjudicate

pytestmark = pytest.mark.gpu

GOLDEN = np.load(os.path.join(os.path.dirname(__file__), "golden", "rvq_golden.npz"))
TIE_CASES = {"exact_rows_and_midpoints"}
ALGOS = ["exact", "auto"]


def make_module(embed, D, C, n_q=None, device="cuda"):
    m = ResidualVectorQuantizer(dimension=D, n_q=len(embed), bins=C, kmeans_init=False)
    sd = m.state_dict()
    for i, e in enumerate(embed):
        sd[f"vq.layers.{i}._codebook.embed"] = torch.from_numpy(np.ascontiguousarray(e))
        sd[f"vq.layers.{i}._codebook.embed_avg"] = torch.from_numpy(np.ascontiguousarray(e))
    m.load_state_dict(sd, strict=True)
    m = m.to(device).eval()
    if n_q is not None:
        m.set_num_codebooks(n_q)
    return m


def to_dev(x_np):
    t = torch.from_numpy(x_np)  # keeps numpy's strides, also for the non-contiguous style-conditioner view
    d = t.to("cuda")
    if t.numel() and not t.is_contiguous():
        assert d.stride() == t.stride()
    return d


def same_bits(a: np.ndarray, b: np.ndarray) -> bool:
    """Bitwise equality (signed zeros included); NaN matches NaN whatever its payload."""
    if a.shape != b.shape:
        return False
    na, nb = np.isnan(a), np.isnan(b)
    if not np.array_equal(na, nb):
        return False
    return np.array_equal(a.view(np.uint32)[~na], b.view(np.uint32)[~nb])


@pytest.mark.parametrize("algo", ALGOS)
@pytest.mark.parametrize("name", list(gen.CASES))
def test_golden_cases(name, algo):
    case = gen.build_case(name)
    cfg = case["cfg"]
    m = make_module(case["embed"], cfg["D"], cfg["C"], case["n_q"])
    m.vq.encode_algo = algo
    x = to_dev(case["x"])
    want_ref = GOLDEN[f"{name}/codes"].astype(np.int64)

    with torch.no_grad():
        codes = m.encode(x)
        res = m(x, 50)
        dec = m.decode(codes)
    B, K, T = want_ref.shape
    assert codes.dtype == torch.int64 and tuple(codes.shape) == (B, K, T)
    if codes.numel():
        assert all(s == w for s, w, n in zip(codes.stride(), GOLDEN[f"{name}/codes_strides"], codes.shape) if n > 1)

    # 1. bit exact against the oracle (same canonical arithmetic)
    o_q, o_codes = oracle.forward(case["x"], case["embed"], case["n_q"])
    got = codes.cpu().numpy()
    assert np.array_equal(got, o_codes), f"{(got != o_codes).sum()} codes differ from the oracle"

    # 2. against the unmodified reference
    n_equal, near, wrong, n_cmp = adjudicate(case["x"], case["embed"], got, want_ref)
    assert wrong == 0
    assert near <= (T // 2 if name in TIE_CASES else 1)

    # 3. forward == (decode o encode), bit for bit; shapes, strides, bandwidth, penalty, metrics
    assert isinstance(res, QuantizedResult)
    assert torch.equal(res.codes, codes)
    assert tuple(res.x.shape) == (B, cfg["D"], T) and res.x.dtype == torch.float32
    rx, dx = res.x.cpu().numpy(), dec.cpu().numpy()
    assert same_bits(np.ascontiguousarray(rx), np.ascontiguousarray(dx))
    assert same_bits(np.ascontiguousarray(rx), np.ascontiguousarray(o_q))
    if res.x.numel():
        ws = tuple(GOLDEN[f"{name}/quant_strides"])
        assert all(s == w for s, w, n in zip(res.x.stride(), ws, res.x.shape) if n > 1)
        assert all(s == w for s, w, n in zip(dec.stride(), ws, dec.shape) if n > 1)
    assert res.bandwidth.dim() == 0 and res.bandwidth.device == x.device and res.bandwidth.dtype == x.dtype
    assert float(res.bandwidth) == pytest.approx(float(GOLDEN[f"{name}/bandwidth"]), rel=1e-6)
    assert res.penalty.dim() == 0 and float(res.penalty) == float(GOLDEN[f"{name}/penalty"]) == 0.0
    assert res.metrics == {}
    assert not res.x.requires_grad

    # 4. decode of the REFERENCE's codes reproduces the reference's output hash
    with torch.no_grad():
        dref = m.decode(torch.from_numpy(want_ref).cuda())
    phys = np.ascontiguousarray(dref.cpu().numpy().transpose(0, 2, 1))
    assert gen.sha(phys) == str(GOLDEN[f"{name}/quant_sha"])


def test_decode_only_cases_int32_fewer_stages_noncontiguous_and_errors():
    embed = gen.codebooks(41, 4, 2048, 128)
    m = make_module(embed, 128, 2048)
    codes_np = gen.randint(42, (3, 4, 500), 2048)
    codes = torch.from_numpy(codes_np).cuda()
    with torch.no_grad():
        q = m.decode(codes)
        q32 = m.decode(codes.to(torch.int32))
        q2 = m.decode(codes[:, :2])
        kbt = codes.transpose(0, 1).contiguous()          # stored [K, B, T] like encode's output
        qv = m.decode(kbt.transpose(0, 1))                 # strided [B, K, T] view
    phys = np.ascontiguousarray(q.cpu().numpy().transpose(0, 2, 1))
    assert gen.sha(phys) == str(GOLDEN["decode_c4/quant_sha"])
    assert q.stride() == (500 * 128, 1, 128)
    assert torch.equal(q, q32) and torch.equal(q, qv)
    phys2 = np.ascontiguousarray(q2.cpu().numpy().transpose(0, 2, 1))
    assert gen.sha(phys2) == str(GOLDEN["decode_c4/quant_k2_sha"])
    assert same_bits(phys, np.ascontiguousarray(oracle.decode(codes_np, embed).transpose(0, 2, 1)))

    for bad_value in (2048, -1, 1 << 40):
        bad = codes.clone()
        bad[1, 2, 3] = bad_value
        with pytest.raises(IndexError):
            m.decode(bad)
    # the error word is cleared: the next valid call works
    with torch.no_grad():
        assert torch.equal(m.decode(codes), q)
    with pytest.raises(IndexError):
        m.decode(torch.zeros(1, 5, 4, dtype=torch.long, device="cuda"))  # more streams than layers
    with pytest.raises(RuntimeError):
        m.decode(codes.to(torch.int16))
    with pytest.raises(RuntimeError):
        m.encode(torch.randn(1, 128, 8, device="cuda", dtype=torch.float64))
    with pytest.raises(RuntimeError):
        m.encode(torch.randn(1, 128, 8, device="cuda", dtype=torch.float16))


def test_negative_zero_rows_decode_to_positive_zero():
    e = [np.full((8, 16), -0.0, dtype=np.float32)]
    m = make_module(e, 16, 8)
    with torch.no_grad():
        q = m.decode(torch.zeros(1, 1, 4, dtype=torch.long, device="cuda"))
        r = m(torch.zeros(1, 16, 4, device="cuda"), 50)
    assert not np.signbit(q.cpu().numpy()).any()
    assert not np.signbit(r.x.cpu().numpy()).any()


@pytest.mark.parametrize("algo", ALGOS)
def test_c2_slice_against_oracle(algo):
    """MusicGen shape, 8 x 1500 frames: bit exact against the oracle on all 48 000 codes."""
    embed = gen.codebooks(7, 4, 2048, 128)
    x_np = gen.normalish(77, (8, 128, 1500))
    m = make_module(embed, 128, 2048)
    m.vq.encode_algo = algo
    with torch.no_grad():
        res = m(torch.from_numpy(x_np).cuda(), 50)
    o_q, o_codes = oracle.forward(x_np, embed)
    assert np.array_equal(res.codes.cpu().numpy(), o_codes)
    assert same_bits(np.ascontiguousarray(res.x.cpu().numpy()), np.ascontiguousarray(o_q))


@pytest.mark.parametrize("algo", ALGOS)
def test_c3_deep_residual_against_oracle(algo):
    embed = gen.codebooks(8, 32, 1024, 128)
    x_np = gen.normalish(78, (2, 128, 750))
    m = make_module(embed, 128, 1024)
    m.vq.encode_algo = algo
    with torch.no_grad():
        codes = m.encode(torch.from_numpy(x_np).cuda())
    assert np.array_equal(codes.cpu().numpy(), oracle.encode(x_np, embed))


@pytest.mark.parametrize("algo", ALGOS)
def test_full_size_c2_properties(algo):
    """BASELINE config 2 at full size (64 x 1500 frames): size-independent properties.
    - frame independence: encoding a batch equals encoding its pieces (any split of B*T is valid)
    - forward(x).x == decode(encode(x)) bit for bit
    - a 1-in-16 sample of the batch items equals the oracle."""
    torch.manual_seed(0)
    m = ResidualVectorQuantizer(dimension=128, n_q=4, bins=2048, kmeans_init=False).cuda().eval()
    m.vq.encode_algo = algo
    g = torch.Generator(device="cuda").manual_seed(1234)
    x = torch.randn(64, 128, 1500, device="cuda", generator=g)
    with torch.no_grad():
        res = m(x, 50)
        codes = m.encode(x)
        dec = m.decode(codes)
        halves = torch.cat([m.encode(x[:32]), m.encode(x[32:])], dim=0)
        time_split = torch.cat([m.encode(x[:, :, :700].contiguous()), m.encode(x[:, :, 700:])], dim=2)
    assert torch.equal(res.codes, codes) and torch.equal(codes, halves) and torch.equal(codes, time_split)
    assert torch.equal(res.x, dec) and res.x.stride() == (1500 * 128, 1, 128)
    assert int(codes.min()) >= 0 and int(codes.max()) < 2048
    embed = [l._codebook.embed.cpu().numpy() for l in m.vq.layers]
    xs = x[::16, :, ::1].cpu().numpy()
    assert np.array_equal(codes[::16].cpu().numpy(), oracle.encode(xs, embed))
    # fewer active stages = a prefix of the codes (the chain is greedy)
    with torch.no_grad():
        m.set_num_codebooks(2)
        assert torch.equal(m.encode(x), codes[:, :2])


def test_full_size_c4_decode_properties():
    """BASELINE config 4 (B=256 x T=3000, K=4): decode equals the sum of per-stage decodes taken in
    order, and a strided sample equals the oracle."""
    torch.manual_seed(0)
    m = ResidualVectorQuantizer(dimension=128, n_q=4, bins=2048, kmeans_init=False).cuda().eval()
    g = torch.Generator(device="cuda").manual_seed(1234)
    codes = torch.randint(0, 2048, (256, 4, 3000), device="cuda", generator=g)
    with torch.no_grad():
        q = m.decode(codes)
        acc = torch.zeros_like(q)
        for k in range(4):
            acc = acc + m.vq.layers[k].decode(codes[:, k])
    assert q.shape == (256, 128, 3000) and q.stride() == (3000 * 128, 1, 128)
    assert torch.equal(q, acc)
    embed = [l._codebook.embed.cpu().numpy() for l in m.vq.layers]
    sample = codes[::37].cpu().numpy()
    want = oracle.decode(sample, embed)
    assert same_bits(np.ascontiguousarray(q[::37].cpu().numpy()), np.ascontiguousarray(want))


def test_pack_follows_codebook_updates_and_devices():
    m = ResidualVectorQuantizer(dimension=32, n_q=2, bins=64, kmeans_init=False).cuda().eval()
    x = torch.randn(2, 32, 50, device="cuda")
    with torch.no_grad():
        c0 = m.encode(x)
        m.vq.layers[0]._codebook.embed.mul_(-1.0)         # in-place update bumps the version counter
        c1 = m.encode(x)
        embed = [l._codebook.embed.cpu().numpy() for l in m.vq.layers]
        assert np.array_equal(c1.cpu().numpy(), oracle.encode(x.cpu().numpy(), embed))
        sd = {k: v.clone() for k, v in m.state_dict().items()}
        sd["vq.layers.1._codebook.embed"] = torch.randn(64, 32)
        m.load_state_dict(sd)                               # strict load, copy_ in place
        c2 = m.encode(x)
        embed = [l._codebook.embed.cpu().numpy() for l in m.vq.layers]
        assert np.array_equal(c2.cpu().numpy(), oracle.encode(x.cpu().numpy(), embed))
    assert not torch.equal(c0, c1)


def test_runs_on_the_callers_stream():
    m = ResidualVectorQuantizer(dimension=128, n_q=4, bins=2048, kmeans_init=False).cuda().eval()
    x = torch.randn(4, 128, 333, device="cuda")
    with torch.no_grad():
        want = m.encode(x)
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            got = m.encode(x)
            dec = m.decode(got)
        s.synchronize()
    assert torch.equal(got, want)
    assert torch.equal(dec, m.decode(want))


def test_per_layer_api_matches_single_stage_oracle():
    embed = gen.codebooks(5, 2, 256, 64)
    m = make_module(embed, 64, 256)
    x_np = gen.normalish(55, (2, 64, 40))
    x = torch.from_numpy(x_np).cuda()
    with torch.no_grad():
        idx = m.vq.layers[0].encode(x)                     # [B, T]
        q = m.vq.layers[0].decode(idx)                     # [B, D, T]
    want = oracle.encode(x_np, embed[:1])[:, 0]
    assert np.array_equal(idx.cpu().numpy(), want)
    assert same_bits(np.ascontiguousarray(q.cpu().numpy()),
                     np.ascontiguousarray(oracle.decode(want[:, None], embed[:1])))


def test_abi_rejects_bad_arguments():
    from audiocraft_b200 import _lib
    lib = _lib.load()
    st = _lib.strides3((1, 1, 1))
    assert lib.rvq_encode_f32(None, st, 1, 128, 1, None, 4, 2048, 4, None, None, None, 0, 0, None) == _lib.RVQ_EINVAL
    assert "null" in _lib.last_error()
    assert lib.rvq_encode_f32(None, st, 1, 128, 1, None, 4, 2048, 9, None, None, None, 0, 0, None) == _lib.RVQ_EINVAL
    assert lib.rvq_encode_f32(None, st, 1, 130, 1, None, 4, 2048, 4, None, None, None, 0, 0, None) != 0
    assert lib.rvq_decode_f32(None, 2, st, 1, 1, 1, None, 4, 2048, 128, None, st, None, None) == _lib.RVQ_EINVAL
    assert kernels.launch_count() >= 0


def test_full_size_c3_deep_residual_properties():
    """BASELINE config 3 at full size (K=32, C=1024, 64 x 2250 frames): prefix property of the greedy chain,
    forward == decode(encode), and a 1-in-32 sample of the batch items against the oracle."""
    torch.manual_seed(0)
    m = ResidualVectorQuantizer(dimension=128, n_q=32, bins=1024, kmeans_init=False).cuda().eval()
    g = torch.Generator(device="cuda").manual_seed(1234)
    x = torch.randn(64, 128, 2250, device="cuda", generator=g)
    with torch.no_grad():
        codes = m.encode(x)
        res = m(x, 75)
        m.set_num_codebooks(8)
        codes8 = m.encode(x)
        m.set_num_codebooks(32)
    assert codes.shape == (64, 32, 2250) and torch.equal(res.codes, codes)
    assert torch.equal(codes8, codes[:, :8])
    with torch.no_grad():
        assert torch.equal(res.x, m.decode(codes))
    embed = [l._codebook.embed.cpu().numpy() for l in m.vq.layers]
    assert np.array_equal(codes[::32].cpu().numpy(), oracle.encode(x[::32].cpu().numpy(), embed))


def test_c5_shard_size_properties():
    """One rank's share of BASELINE config 5 at 8 GPUs (K=8, 2^21 frames): encoding the shard equals encoding its
    pieces; exact and tensor-core paths agree on a slice; sampled oracle check."""
    torch.manual_seed(0)
    m = ResidualVectorQuantizer(dimension=128, n_q=8, bins=2048, kmeans_init=False).cuda().eval()
    g = torch.Generator(device="cuda").manual_seed(1234 + 3)
    x = torch.randn(1024, 128, 2048, device="cuda", generator=g)
    with torch.no_grad():
        codes = m.encode(x)
        parts = torch.cat([m.encode(x[i:i + 256]) for i in range(0, 1024, 256)], dim=0)
        assert torch.equal(codes, parts)
        m.vq.encode_algo = "exact"
        assert torch.equal(m.encode(x[:16]), codes[:16])
        m.vq.encode_algo = "auto"
        q = m.decode(codes[:64])
    assert q.shape == (64, 128, 2048)
    embed = [l._codebook.embed.cpu().numpy() for l in m.vq.layers]
    sl = x[5:6, :, :512].cpu().numpy()
    assert np.array_equal(codes[5:6, :, :512].cpu().numpy(), oracle.encode(sl, embed))


@pytest.mark.parametrize("name,K,C,B,T,fr", [("c2", 4, 2048, 64, 1500, 50), ("c3", 32, 1024, 64, 2250, 75),
                                              ("c5_shard", 8, 2048, 1024, 2048, 50)])
def test_full_size_exact_scan_equals_tensor_core_search(name, K, C, B, T, fr):
    """Every code of the full-size batches of BASELINE configs 2, 3 and 5 (one rank's share at 8 GPUs): the
    certified tensor-core search and the CUDA-core scan that scores every code in the canonical order emit the
    same codes, bit for bit (the scan is what the oracle pins; the sampled oracle checks are in the tests above)."""
    torch.manual_seed(0)
    m = ResidualVectorQuantizer(dimension=128, n_q=K, bins=C, kmeans_init=False).cuda().eval()
    g = torch.Generator(device="cuda").manual_seed(1234)
    x = torch.randn(B, 128, T, device="cuda", generator=g)
    with torch.no_grad():
        m.vq.encode_algo = "tc"
        tc = m.encode(x)
        m.vq.encode_algo = "exact"
        ex = m.encode(x)
    diff = int((tc != ex).sum())
    print(f"{name}: {tc.numel()} codes, {diff} differ between the exact scan and the tensor-core search")
    assert diff == 0


def test_full_c2_batch_against_the_reference_codes():
    """All 384 000 codes of BASELINE config 2 against the codes of the unmodified reference
    (tests/golden/rvq_golden_c2_full.npz): equal except fp32 near-ties, which are counted, adjudicated in fp64
    (tests/parity.py) and bounded.  SURVEY 8(c): "report the count (target 0)"."""
    from tests.golden import make_golden_c2_full as full
    gold = np.load(os.path.join(os.path.dirname(__file__), "golden", "rvq_golden_c2_full.npz"))
    x_np, embed = full.inputs()
    assert gen.sha(x_np) == str(gold["x_sha"]) and gen.sha(np.stack(embed)) == str(gold["embed_sha"])
    want = gold["codes"].astype(np.int64)
    m = make_module(embed, 128, 2048)
    with torch.no_grad():
        x = torch.from_numpy(x_np).cuda()
        got = {}
        for algo in ALGOS:
            m.vq.encode_algo = algo
            got[algo] = m.encode(x)
        dec = m.decode(torch.from_numpy(want).cuda())
    assert torch.equal(got["exact"], got["auto"])
    n_equal, near, wrong, n_cmp = adjudicate(x_np, embed, got["auto"].cpu().numpy(), want)
    print(f"full c2 vs reference: {want.size} codes, {want.size - n_equal} differ, {near} adjudicated fp32 near-ties, "
          f"{wrong} wrong, {n_cmp} compared")
    assert wrong == 0
    assert near <= 8, f"{near} near-ties in 96 000 frames"
    # decode of the reference's codes: bit exact against the reference's own dequantized output
    assert gen.sha(np.ascontiguousarray(dec.cpu().numpy().transpose(0, 2, 1))) == str(gold["quant_sha"])
