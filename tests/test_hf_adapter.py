"""HuggingFace EnCodec's RVQ twin: same state, same results as
transformers.models.encodec.modeling_encodec.EncodecResidualVectorQuantizer, built here from a config with random
codebooks (no checkpoint download)."""
import numpy as np
import pytest
import torch

import oracle
from audiocraft_b200.hf_adapter import HFEncodecQuantizer, patch_hf_model
from tests.golden import gen
from tests.parity import adjudicate

me = pytest.importorskip("transformers.models.encodec.modeling_encodec")
from transformers import EncodecConfig  # noqa: E402


def hf_quantizer(num_codebooks=8, codebook_size=1024, dim=128, seed=5):
    cfg = EncodecConfig(codebook_size=codebook_size, codebook_dim=dim, hidden_size=dim)
    q = me.EncodecResidualVectorQuantizer(cfg)
    q.layers = q.layers[:num_codebooks]
    q.num_quantizers = num_codebooks
    for i, e in enumerate(gen.codebooks(seed, num_codebooks, codebook_size, dim)):
        q.layers[i].codebook.embed.copy_(torch.from_numpy(e))
        q.layers[i].codebook.embed_avg.copy_(torch.from_numpy(e))
    return q.eval(), cfg


def test_state_mapping_round_trip_and_bandwidths():
    hf, cfg = hf_quantizer()
    mine = HFEncodecQuantizer.from_hf(hf)
    sd_hf, sd_mine = hf.state_dict(), mine.state_dict()
    assert list(sd_hf.keys()) == list(sd_mine.keys())
    for k in sd_hf:
        assert torch.equal(sd_hf[k], sd_mine[k]), k
    hf2, _ = hf_quantizer(seed=6)
    hf2.load_state_dict(mine.state_dict(), strict=True)       # and back into transformers' class
    assert torch.equal(hf2.layers[3].codebook.embed, hf.layers[3].codebook.embed)
    for bw in (None, 0.0, 1.5, 3.0, 6.0, 24.0):
        assert mine.get_num_quantizers_for_bandwidth(bw) == hf.get_num_quantizers_for_bandwidth(bw)
    with pytest.raises(RuntimeError):                          # no CPU fallback
        mine.encode(torch.zeros(1, 128, 4))


def test_oracle_equals_the_hf_twin_on_cpu():
    hf, _ = hf_quantizer(num_codebooks=4)
    x = gen.normalish(91, (2, 128, 300))
    with torch.no_grad():
        codes = hf.encode(torch.from_numpy(x), None)                        # [K, B, T]
        q = hf.decode(codes)
    embed = [l.codebook.embed.numpy() for l in hf.layers]
    got = oracle.encode(x, embed)
    _, near, wrong, _ = adjudicate(x, embed, got, codes.transpose(0, 1).numpy())
    assert wrong == 0 and near <= 1
    assert np.array_equal(oracle.decode(codes.transpose(0, 1).numpy(), embed), q.numpy())


@pytest.mark.gpu
def test_adapter_on_the_kernels_equals_the_hf_twin():
    hf, _ = hf_quantizer(num_codebooks=8)
    mine = HFEncodecQuantizer.from_hf(hf).cuda()
    x = gen.normalish(92, (3, 128, 500))
    embed = [l.codebook.embed.numpy() for l in hf.layers]
    with torch.no_grad():
        for bw in (None, 3.0):
            want = hf.encode(torch.from_numpy(x), bw)
            got = mine.encode(torch.from_numpy(x).cuda(), bw)
            assert got.shape == want.shape and got.dtype == want.dtype
            _, near, wrong, _ = adjudicate(x, embed, got.transpose(0, 1).cpu().numpy(), want.transpose(0, 1).numpy())
            assert wrong == 0 and near <= 1
        q_want = hf.decode(want)
        q_got = mine.decode(want.cuda())
    assert q_got.shape == q_want.shape and torch.equal(q_got.cpu(), q_want)

    class Holder(torch.nn.Module):  # what HFEncodecCompressionModel.decode_latent touches (encodec.py:360-362)
        def __init__(self, quantizer):
            super().__init__()
            self.quantizer = quantizer
    model = patch_hf_model(Holder(hf)).cuda()
    codes_bkt = want.transpose(0, 1).cuda()
    assert torch.equal(model.quantizer.decode(codes_bkt.transpose(0, 1)).cpu(), q_want)
