"""CPU checks of the drop-in boundary: constructor, state_dict contract, C-ABI symbols, and that
the product refuses to run without CUDA (no CPU fallback)."""
import ctypes
import os
import re

import pytest
import torch

import audiocraft_b200
from audiocraft_b200 import _lib
from audiocraft_b200.quantization import (BaseQuantizer, DummyQuantizer, QuantizedResult,
                                          ResidualVectorQuantizer)
from oracle import ref_shim

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_loads_and_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, "include", "rvq_b200.h")).read()
    declared = set(re.findall(r"\b(rvq_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations found"
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/rvq_b200.h but not exported"
    assert declared == set(_lib.SIGNATURES), "ctypes binding and header disagree"
    assert _lib.load().rvq_b200_abi_version() == _lib.ABI_VERSION == 2


def test_pack_bytes_is_host_only_and_validates():
    lib = _lib.load()
    assert lib.rvq_pack_bytes(4, 2048, 128) > 4 * 2048 * 128 * 4
    assert lib.rvq_pack_bytes(0, 2048, 128) < 0
    assert lib.rvq_pack_bytes(65, 2048, 128) < 0


def test_constructor_defaults_and_state_dict_keys():
    m = ResidualVectorQuantizer()
    assert (m.dimension, m.n_q, m.bins, m.decay, m.kmeans_init) == (256, 8, 1024, 0.99, True)
    assert isinstance(m, BaseQuantizer) and isinstance(m, torch.nn.Module)
    m = ResidualVectorQuantizer(dimension=128, n_q=4, bins=2048, kmeans_init=False)
    sd = m.state_dict()
    want = {f"vq.layers.{i}._codebook.{b}" for i in range(4) for b in ("inited", "cluster_size", "embed", "embed_avg")}
    assert set(sd) == want
    assert sd["vq.layers.0._codebook.embed"].shape == (2048, 128)
    assert sd["vq.layers.0._codebook.embed_avg"].shape == (2048, 128)
    assert sd["vq.layers.0._codebook.cluster_size"].shape == (2048,)
    assert sd["vq.layers.0._codebook.inited"].shape == (1,) and float(sd["vq.layers.0._codebook.inited"]) == 1.0
    assert all(v.dtype == torch.float32 for v in sd.values())
    assert len(list(m.parameters())) == 0 and len(list(m.buffers())) == 16
    bound = (6.0 / 128) ** 0.5
    e = sd["vq.layers.0._codebook.embed"]
    assert float(e.abs().max()) <= bound and float(e.abs().max()) > 0.9 * bound
    zeros = ResidualVectorQuantizer(dimension=16, n_q=2, bins=8, kmeans_init=True).state_dict()
    assert float(zeros["vq.layers.0._codebook.embed"].abs().max()) == 0.0
    assert float(zeros["vq.layers.0._codebook.inited"]) == 0.0


def test_positional_construction_like_the_conditioners():
    # conditioners.py:928-929 and :1037
    m = ResidualVectorQuantizer(512, n_q=6, q_dropout=False, bins=1024, threshold_ema_dead_code=2)
    assert m.dimension == 512 and m.total_codebooks == 6 and m.num_codebooks == 6
    m.set_num_codebooks(3)
    assert m.num_codebooks == 3 and m.total_codebooks == 6
    with pytest.raises(AssertionError):
        m.set_num_codebooks(0)
    with pytest.raises(AssertionError):
        m.set_num_codebooks(7)


@pytest.mark.skipif(not ref_shim.available(), reason="reference tree not present")
def test_state_dict_round_trip_with_the_reference():
    q = ref_shim.load()
    torch.manual_seed(0)
    ref = q.ResidualVectorQuantizer(dimension=32, n_q=3, bins=50, kmeans_init=False)
    mine = ResidualVectorQuantizer(dimension=32, n_q=3, bins=50, kmeans_init=False)
    assert list(ref.state_dict()) == list(mine.state_dict())
    mine.load_state_dict(ref.state_dict(), strict=True)
    ref.load_state_dict(mine.state_dict(), strict=True)
    for (ka, va), (kb, vb) in zip(ref.state_dict().items(), mine.state_dict().items()):
        assert ka == kb and torch.equal(va, vb)
    import inspect
    ref_params = inspect.signature(q.ResidualVectorQuantizer.__init__).parameters
    my_params = inspect.signature(ResidualVectorQuantizer.__init__).parameters
    assert [(n, p.default) for n, p in ref_params.items()] == [(n, p.default) for n, p in my_params.items()]


def test_no_cpu_fallback():
    m = ResidualVectorQuantizer(dimension=16, n_q=2, bins=8, kmeans_init=False).eval()
    x = torch.randn(1, 16, 10)
    with pytest.raises(RuntimeError, match="CUDA"):
        m.encode(x)
    with pytest.raises(RuntimeError, match="CUDA"):
        m(x, 50)
    with pytest.raises(RuntimeError, match="CUDA"):
        m.decode(torch.zeros(1, 2, 10, dtype=torch.long))


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "audiocraft_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), f
                assert "rvq_oracle" not in src, f


def test_dummy_quantizer_surface():
    d = DummyQuantizer()
    x = torch.randn(2, 4, 6)
    res = d(x, 50)
    assert isinstance(res, QuantizedResult) and res.codes.shape == (2, 1, 4, 6) and res.metrics == {}
    assert torch.equal(d.decode(d.encode(x)), x)
    assert d.total_codebooks == 1 and d.num_codebooks == 1
    with pytest.raises(AttributeError):
        d.set_num_codebooks(2)
    assert float(res.bandwidth) == pytest.approx(x.unsqueeze(1).numel() * 32 * 50 / 1000 / 2)


def test_package_exports():
    assert audiocraft_b200.ResidualVectorQuantizer is ResidualVectorQuantizer
