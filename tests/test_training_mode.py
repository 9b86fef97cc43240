"""Training-mode branch of the quantizer: k-means init, EMA codebook updates,
dead-code expiry, commitment loss, straight-through gradients, q_dropout.

CPU part: the host logic against the unmodified reference under the same seeds.  The nearest-code search of
the product is CUDA-only, so here it is replaced by a test-side torch search (test infrastructure, never
shipped).  GPU part: the same logic end to end on the kernels, checked through its invariants.
"""
import numpy as np
import pytest
import torch

from audiocraft_b200.quantization import ResidualVectorQuantizer, core_vq
from oracle import ref_shim


def _torch_search(self, flat):
    e = self.embed.t()
    dist = -(flat.pow(2).sum(1, keepdim=True) - 2 * flat @ e + e.pow(2).sum(0, keepdim=True))
    return dist.max(dim=-1).indices


CONFIGS = [
    dict(n_q=8, dimension=16, bins=8),                                     # tests/quantization/test_vq.py:16
    dict(n_q=4, dimension=32, bins=400, q_dropout=True),                   # builders.py:277 + q_dropout
    dict(n_q=3, dimension=32, bins=64, kmeans_init=False, orthogonal_reg_weight=0.5,
         orthogonal_reg_active_codes_only=True, orthogonal_reg_max_codes=16),
]


@pytest.mark.skipif(not ref_shim.available(), reason="reference tree not present")
@pytest.mark.parametrize("kw", CONFIGS)
def test_training_steps_match_reference_on_cpu(kw, monkeypatch):
    monkeypatch.setattr(core_vq.EuclideanCodebook, "_search", _torch_search)
    q = ref_shim.load()
    torch.manual_seed(3)
    ref = q.ResidualVectorQuantizer(**kw)
    mine = ResidualVectorQuantizer(**kw)
    mine.load_state_dict(ref.state_dict())
    x = torch.randn(2, kw["dimension"], 300)
    for step in range(3):
        outs = []
        for m in (ref, mine):
            torch.manual_seed(100 + step)                      # same random draws (k-means, expiry, dropout)
            xi = x.clone().requires_grad_(True)
            res = m(xi, 50)
            res.x.sum().backward()
            outs.append((res, xi.grad))
        (rr, gr), (rm, gm) = outs
        assert torch.equal(rr.codes, rm.codes)
        assert torch.allclose(rr.x, rm.x, atol=1e-6)
        assert float(rr.penalty.detach()) == pytest.approx(float(rm.penalty.detach()), rel=1e-5)
        assert float(rr.bandwidth) == float(rm.bandwidth)
        assert torch.allclose(gr, gm)
    for (ka, a), (kb, b) in zip(ref.state_dict().items(), mine.state_dict().items()):
        assert ka == kb and torch.allclose(a, b, atol=1e-5, rtol=1e-4), ka
    # (eval-mode use after training runs on the kernels: covered by the GPU tests below)


@pytest.mark.gpu
def test_reference_test_rvq_on_gpu():
    """tests/quantization/test_vq.py:14-18 of the reference, on the B200 path (default = training mode,
    k-means initialisation on the first batch)."""
    torch.manual_seed(0)
    x = torch.randn(1, 16, 2048, device="cuda")
    vq = ResidualVectorQuantizer(n_q=8, dimension=16, bins=8).cuda()
    res = vq(x, 1.)
    assert res.x.shape == torch.Size([1, 16, 2048])
    assert res.codes.shape == (1, 8, 2048) and int(res.codes.min()) >= 0 and int(res.codes.max()) < 8
    assert all(float(l._codebook.inited) == 1.0 for l in vq.vq.layers)
    assert float(res.penalty.detach()) > 0


@pytest.mark.gpu
def test_training_step_invariants_on_gpu():
    torch.manual_seed(1)
    D, C, K = 32, 400, 4
    vq = ResidualVectorQuantizer(dimension=D, n_q=K, bins=C, kmeans_init=False, decay=0.9).cuda()
    before = {k: v.clone() for k, v in vq.state_dict().items()}
    x = torch.randn(8, D, 128, device="cuda", requires_grad=True)
    res = vq(x, 50)
    res.x.sum().backward()
    assert torch.equal(x.grad, torch.ones_like(x))            # straight-through over the whole chain
    assert res.penalty.requires_grad
    # stage 0 statistics follow from the returned codes: EMA of the histogram and of the per-code sums
    codes0 = res.codes[:, 0].reshape(-1)
    flat = x.detach().transpose(1, 2).reshape(-1, D)
    counts = torch.bincount(codes0, minlength=C).float()
    want_cs = 0.9 * before["vq.layers.0._codebook.cluster_size"] + 0.1 * counts
    # dead codes (cluster_size < 2 at construction: all of them) were re-drawn from the batch first, so
    # only the usage statistics can be predicted exactly here
    assert torch.allclose(vq.state_dict()["vq.layers.0._codebook.cluster_size"], want_cs, atol=1e-5)
    sums = torch.zeros(C, D, device="cuda").index_add_(0, codes0, flat)
    want_avg = 0.9 * before["vq.layers.0._codebook.embed_avg"] + 0.1 * sums
    assert torch.allclose(vq.state_dict()["vq.layers.0._codebook.embed_avg"], want_avg, atol=1e-4)
    # quantized output = sum of the selected rows of the codebooks used for the search (before the update)
    assert res.x.shape == x.shape and torch.isfinite(res.x).all()


@pytest.mark.gpu
def test_debug_model_flow_and_uninitialised_eval_on_gpu():
    """builders.get_debug_compression_model (builders.py:277-279): one training-mode forward to initialise
    the codebooks by k-means, then eval use; and eval forward of an uninitialised module initialises it."""
    torch.manual_seed(2)
    vq = ResidualVectorQuantizer(dimension=32, n_q=4, bins=400).cuda()
    vq(torch.randn(8, 32, 128, device="cuda"), 50)
    vq.eval()
    x = torch.randn(2, 32, 64, device="cuda")
    with torch.no_grad():
        res = vq(x, 50)
        assert torch.equal(res.codes, vq.encode(x)) and torch.equal(res.x, vq.decode(res.codes))
    fresh = ResidualVectorQuantizer(dimension=32, n_q=2, bins=16).cuda().eval()
    assert float(fresh.vq.layers[0]._codebook.inited) == 0.0
    with torch.no_grad():
        r2 = fresh(torch.randn(4, 32, 100, device="cuda"), 50)
    assert float(fresh.vq.layers[0]._codebook.inited) == 1.0 and float(r2.penalty) == 0.0
    with torch.no_grad():
        assert torch.equal(fresh(x, 50).codes, fresh.encode(x))


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["tc_d128_c256", "small_d32_c64"])
def test_training_steps_on_the_kernels_match_the_reference_fixtures(name):
    """Three training steps on the GPU (search on the encode kernels, statistics + EMA + re-normalisation on
    rvq_ema_update_f32) against tests/golden/rvq_golden_training.npz, which the unmodified reference produced on the
    same inputs: codes exactly, the dequantized output and every buffer within 1e-5."""
    import os
    from tests.golden import make_golden_training as mg
    gold = np.load(os.path.join(os.path.dirname(__file__), "golden", "rvq_golden_training.npz"))
    cfg = mg.CONFIGS[name]
    embed, xs = mg.inputs(cfg)
    m = ResidualVectorQuantizer(**mg.kwargs(cfg))
    mg.load_codebooks(m, embed)
    m = m.cuda().train()
    for s in range(mg.STEPS):
        res = m(torch.from_numpy(xs[s]).cuda(), 50)
        want = gold[f"{name}/codes{s}"].astype(np.int64)
        got = res.codes.cpu().numpy()
        assert np.array_equal(got, want), f"step {s}: {(got != want).sum()} of {want.size} codes differ"
        assert float(res.penalty.detach()) == pytest.approx(float(gold[f"{name}/penalty{s}"]), rel=1e-5)
        if cfg["every_step"] or s == mg.STEPS - 1:
            for k, v in m.state_dict().items():
                ref = gold[f"{name}/step{s}/{k}"]
                assert np.allclose(v.cpu().numpy(), ref, rtol=1e-5, atol=1e-5), f"step {s}: buffer {k} differs"
    # eval after training runs on the fused kernels with the trained codebooks
    m.eval()
    with torch.no_grad():
        x = torch.from_numpy(xs[0]).cuda()
        assert torch.equal(m(x, 50).codes, m.encode(x))


@pytest.mark.gpu
def test_ema_statistics_kernel_is_deterministic_and_ordered():
    """rvq_ema_update_f32: counts exact; per-code sums equal the fp32 sum in increasing frame order bit for bit (an
    independent host evaluation of that order), twice in a row; EMA and re-normalisation follow core_vq.py:209-217."""
    from audiocraft_b200 import kernels
    torch.manual_seed(0)
    N, C, D = 5000, 96, 48
    x = torch.randn(N, D, device="cuda")
    codes = torch.randint(0, C - 3, (N,), device="cuda")      # the last three codes stay unused
    decay, eps = 0.9, 1e-5
    bufs0 = [torch.rand(C, device="cuda"), torch.randn(C, D, device="cuda"), torch.zeros(C, D, device="cuda")]
    results = []
    for _ in range(2):
        cs_, ea, em = [b.clone() for b in bufs0]
        counts, sums = kernels.ema_update(x.t().contiguous().t(), codes, cs_, ea, em, decay, eps, want_stats=True)  # strided input
        results.append((counts, sums, cs_, ea, em))
    for a, b in zip(results[0], results[1]):
        assert torch.equal(a, b)
    counts, sums, cs_, ea, em = results[0]
    assert torch.equal(counts, torch.bincount(codes, minlength=C).float())
    xh, ch = x.cpu().numpy(), codes.cpu().numpy()
    want = np.zeros((C, D), dtype=np.float32)
    for n in range(N):                                          # sequential fp32 adds in frame order
        want[ch[n]] = want[ch[n]] + xh[n]
    assert np.array_equal(sums.cpu().numpy(), want)
    cs_ref = bufs0[0] * decay + counts * (1 - decay)
    ea_ref = bufs0[1] * decay + sums * (1 - decay)
    tot = cs_ref.sum()
    em_ref = ea_ref / (((cs_ref + eps) / (tot + C * eps)) * tot).unsqueeze(1)
    assert torch.allclose(cs_, cs_ref, rtol=1e-5) and torch.allclose(ea, ea_ref, rtol=1e-5, atol=1e-6)
    assert torch.allclose(em, em_ref, rtol=1e-5, atol=1e-6)
