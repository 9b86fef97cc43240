"""Pins the CPU oracle (oracle/rvq_oracle.c) against vectors produced by the unmodified reference
(tests/golden/rvq_golden.npz, made by tests/golden/make_golden.py)."""
import os

import numpy as np
import pytest

import oracle
from tests.golden import gen
from tests.parity import adjudicate

GOLDEN = np.load(os.path.join(os.path.dirname(__file__), "golden", "rvq_golden.npz"))
# cases built to contain exact mathematical ties (midpoints between two rows): rounding decides
TIE_CASES = {"exact_rows_and_midpoints"}


@pytest.mark.parametrize("name", list(gen.CASES))
def test_generated_inputs_match_recorded_hashes(name):
    case = gen.build_case(name)
    assert gen.sha(case["x"]) == str(GOLDEN[f"{name}/x_sha"])
    assert gen.sha(np.stack(case["embed"])) == str(GOLDEN[f"{name}/embed_sha"])


@pytest.mark.parametrize("name", list(gen.CASES))
def test_oracle_encode_forward_vs_reference(name):
    case = gen.build_case(name)
    want = GOLDEN[f"{name}/codes"].astype(np.int64)
    quant, codes = oracle.forward(case["x"], case["embed"], case["n_q"])
    assert codes.shape == want.shape and codes.dtype == np.int64
    ws = tuple(GOLDEN[f"{name}/codes_strides"])
    assert all(g // 8 == w for g, w, n in zip(codes.strides, ws, codes.shape) if n > 1) or want.size == 0
    enc = oracle.encode(case["x"], case["embed"], case["n_q"])
    assert np.array_equal(enc, codes)
    n_equal, near, wrong, n_cmp = adjudicate(case["x"], case["embed"], codes, want)
    assert wrong == 0
    if name in TIE_CASES:
        assert near <= want.shape[2] // 2
    else:
        assert near <= 1, f"{near} near-ties"  # c2_slice has one fp32 near-tie
    if near == 0:
        assert np.array_equal(codes, want)
        assert gen.sha(np.ascontiguousarray(quant.transpose(0, 2, 1))) == str(GOLDEN[f"{name}/quant_sha"])


@pytest.mark.parametrize("name", list(gen.CASES))
def test_oracle_decode_of_reference_codes(name):
    """decode() is pure gather + ordered fp32 sum: bit exact, no tolerance."""
    case = gen.build_case(name)
    want = GOLDEN[f"{name}/codes"].astype(np.int64)
    q = oracle.decode(want, case["embed"])
    B, K, T = want.shape
    D = case["cfg"]["D"]
    assert q.shape == (B, D, T) and q.dtype == np.float32
    if want.size:  # strides of size-1 dimensions carry no information
        ws = tuple(GOLDEN[f"{name}/quant_strides"])
        assert all(g // 4 == w for g, w, n in zip(q.strides, ws, q.shape) if n > 1)
    phys = np.ascontiguousarray(q.transpose(0, 2, 1))
    assert gen.sha(phys) == str(GOLDEN[f"{name}/quant_sha"])
    key = f"{name}/quant"
    if key in GOLDEN.files:
        assert np.array_equal(phys, GOLDEN[key], equal_nan=True)
    else:
        assert np.array_equal(phys.reshape(-1)[::97], GOLDEN[f"{name}/quant_sample"], equal_nan=True)


def test_oracle_decode_only_cases():
    embed = gen.codebooks(41, 4, 2048, 128)
    codes = gen.randint(42, (3, 4, 500), 2048)
    q = oracle.decode(codes, embed)
    assert gen.sha(np.ascontiguousarray(q.transpose(0, 2, 1))) == str(GOLDEN["decode_c4/quant_sha"])
    q32 = oracle.decode(codes.astype(np.int32), embed)
    assert np.array_equal(q, q32)
    q2 = oracle.decode(codes[:, :2], embed)  # fewer stages than layers (core_vq.py:401-402)
    assert gen.sha(np.ascontiguousarray(q2.transpose(0, 2, 1))) == str(GOLDEN["decode_c4/quant_k2_sha"])
    bad = codes.copy()
    bad[1, 2, 3] = 2048
    with pytest.raises(IndexError):
        oracle.decode(bad, embed)
    bad[1, 2, 3] = -1
    with pytest.raises(IndexError):
        oracle.decode(bad, embed)


def test_negative_zero_rows_decode_to_positive_zero():
    e = [np.full((8, 16), -0.0, dtype=np.float32)]
    q = oracle.decode(np.zeros((1, 1, 4), dtype=np.int64), e)
    assert np.array_equal(np.signbit(q).astype(np.uint8), GOLDEN["negzero/signbits"])
    assert not np.signbit(q).any()


def test_canonical_dot_definition():
    """The canonical order, restated in numpy: 32 partials, element d -> lane (d%128)//4, fmaf chain,
    then the xor butterfly 16,8,4,2,1."""
    rng = np.random.default_rng(0)
    for D in (4, 16, 32, 100, 128, 256, 512):
        a = rng.standard_normal(D).astype(np.float32)
        b = rng.standard_normal(D).astype(np.float32)
        p = np.zeros(32, dtype=np.float32)
        for d in range(D):
            lane = (d % 128) // 4
            # single-rounded fma via float64 product (exact) + add, then one rounding: exact for fp32 inputs
            p[lane] = np.float32(np.float64(a[d]) * np.float64(b[d]) + np.float64(p[lane]))
        for off in (16, 8, 4, 2, 1):
            p = (p + p[np.arange(32) ^ off]).astype(np.float32)
        assert np.float32(oracle.dot(a, b)) == p[0]


def test_oracle_full_c2_batch_vs_reference():
    """The whole of BASELINE config 2 (96 000 frames, 384 000 codes) against the unmodified reference's codes:
    the count of fp32 near-ties is reported and bounded, nothing else may differ."""
    from tests.golden import make_golden_c2_full as full
    gold = np.load(os.path.join(os.path.dirname(__file__), "golden", "rvq_golden_c2_full.npz"))
    x, embed = full.inputs()
    assert gen.sha(x) == str(gold["x_sha"]) and gen.sha(np.stack(embed)) == str(gold["embed_sha"])
    want = gold["codes"].astype(np.int64)
    codes = oracle.encode(x, embed)
    n_equal, near, wrong, n_cmp = adjudicate(x, embed, codes, want)
    print(f"oracle vs reference on full c2: {want.size - n_equal} of {want.size} codes differ, {near} near-ties, {wrong} wrong")
    assert wrong == 0 and near <= 8
    q = oracle.decode(want, embed)
    assert gen.sha(np.ascontiguousarray(q.transpose(0, 2, 1))) == str(gold["quant_sha"])
