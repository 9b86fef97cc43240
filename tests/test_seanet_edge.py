"""SEANet's edge convolutions in channels-last memory against the reference's StreamableConv1d."""
import importlib
import os
import sys
import types

import pytest
import torch

from audiocraft_b200.seanet_edge import EdgeConv1d

REF = "/root/reference"


def ref_conv():
    path = os.path.join(REF, "audiocraft", "modules", "conv.py")
    if not os.path.isfile(path):
        pytest.skip("reference not available")
    from oracle import ref_shim
    ref_shim.load()
    if "audiocraft.modules" not in sys.modules:
        pkg = types.ModuleType("audiocraft.modules")
        pkg.__path__ = [os.path.join(REF, "audiocraft", "modules")]
        sys.modules["audiocraft.modules"] = pkg
    return importlib.import_module("audiocraft.modules.conv")


@pytest.mark.parametrize("cin,cout,k,norm,causal,T", [(64, 128, 7, "weight_norm", False, 50), (128, 96, 7, "none", True, 33),
                                                      (32, 16, 3, "weight_norm", False, 2), (16, 16, 1, "none", False, 9)])
def test_edge_conv_equals_streamable_conv1d(cin, cout, k, norm, causal, T):
    conv = ref_conv()
    torch.manual_seed(0)
    ref = conv.StreamableConv1d(cin, cout, k, norm=norm, causal=causal, pad_mode="reflect")
    mine = EdgeConv1d.from_reference(ref)
    x = torch.randn(3, cin, T)
    with torch.no_grad():
        want = ref(x)
        got = mine(x)
        got_cl = mine(x.transpose(1, 2).contiguous().transpose(1, 2))      # frame-major input, as the quantizer returns it
    assert got.shape == want.shape and got.stride() == (T * cout, 1, cout)  # [B, T, C] memory
    assert torch.allclose(got, want, atol=1e-5, rtol=1e-5) and torch.allclose(got_cl, want, atol=1e-5, rtol=1e-5)


@pytest.mark.gpu
def test_encoder_edge_feeds_the_encode_kernel_without_a_copy():
    """Last encoder conv -> quantizer -> first decoder conv, all on [B, T, D] memory: same codes as the [B, D, T] path,
    same decoder-side activations (within the conv's fp32 tolerance)."""
    from audiocraft_b200.quantization import ResidualVectorQuantizer
    torch.manual_seed(0)
    enc_last = torch.nn.Conv1d(256, 128, 7)
    dec_first = torch.nn.Conv1d(128, 256, 7)
    e = EdgeConv1d(enc_last.weight, enc_last.bias).cuda()
    d = EdgeConv1d(dec_first.weight, dec_first.bias).cuda()
    m = ResidualVectorQuantizer(dimension=128, n_q=4, bins=2048, kmeans_init=False).cuda().eval()
    h = torch.randn(2, 256, 400, device="cuda")
    with torch.no_grad():
        emb = e(h)                                            # [B, 128, T] view of [B, T, 128]
        assert emb.stride() == (400 * 128, 1, 128)
        codes = m.encode(emb)                                 # frame-major fast path of the kernel
        assert torch.equal(codes, m.encode(emb.contiguous()))
        z = m.decode(codes)
        assert z.stride() == (400 * 128, 1, 128)
        y = d(z)
        want = torch.nn.functional.conv1d(torch.nn.functional.pad(z.contiguous(), (3, 3), mode="reflect"),
                                          dec_first.weight.cuda(), dec_first.bias.cuda())
    assert torch.allclose(y, want, atol=1e-4, rtol=1e-4)
