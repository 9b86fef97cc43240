"""HostPipeline (pinned host in -> pinned host out, chunked over three streams) returns exactly what the
plain device calls return."""
import numpy as np
import pytest
import torch

from audiocraft_b200.host_pipeline import HostPipeline
from audiocraft_b200.quantization import ResidualVectorQuantizer

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("B,chunk", [(10, 4), (3, 16), (16, 16)])
def test_host_pipeline_matches_device_calls(B, chunk):
    torch.manual_seed(0)
    m = ResidualVectorQuantizer(dimension=128, n_q=4, bins=2048, kmeans_init=False).cuda().eval()
    pipe = HostPipeline(m, chunk_items=chunk)
    T = 333
    for call in range(3):  # repeated calls reuse the staging slots
        x = torch.randn(B, 128, T).pin_memory()
        codes_h = torch.full((B, 4, T), -1, dtype=torch.int64).pin_memory()
        quant_h = torch.full((B, T, 128), float("nan")).pin_memory()
        pipe.encode_decode(x, codes_h, quant_h).synchronize()
        with torch.no_grad():
            want_codes = m.encode(x.cuda())
            want_q = m.decode(want_codes)
        assert torch.equal(codes_h, want_codes.cpu())
        assert np.array_equal(quant_h.numpy().view(np.uint32),
                              np.ascontiguousarray(want_q.transpose(1, 2).cpu().numpy()).view(np.uint32))
    assert m.vq.check_codes is True  # restored


def test_host_pipeline_rejects_pageable_or_device_buffers():
    m = ResidualVectorQuantizer(dimension=128, n_q=2, bins=256, kmeans_init=False).cuda().eval()
    pipe = HostPipeline(m)
    x = torch.randn(2, 128, 8)
    with pytest.raises(ValueError):
        pipe.encode_decode(x, torch.empty(2, 2, 8, dtype=torch.int64).pin_memory(), torch.empty(2, 8, 128).pin_memory())


def test_host_pipeline_with_changing_lengths():
    """Variable-length audio: the staging slots grow once and are reused as views; calls are issued without waiting for
    the previous one (the pipeline is asynchronous), and every call must still see its own data."""
    torch.manual_seed(1)
    m = ResidualVectorQuantizer(dimension=128, n_q=4, bins=2048, kmeans_init=False).cuda().eval()
    pipe = HostPipeline(m, chunk_items=4)
    jobs = []
    for T in (200, 700, 150, 701, 64):
        x = torch.randn(6, 128, T).pin_memory()
        codes_h = torch.full((6, 4, T), -1, dtype=torch.int64).pin_memory()
        quant_h = torch.full((6, T, 128), float("nan")).pin_memory()
        jobs.append((x, codes_h, quant_h, pipe.encode_decode(x, codes_h, quant_h)))
    pipe.synchronize()
    for x, codes_h, quant_h, ev in jobs:
        ev.synchronize()
        with torch.no_grad():
            want = m.encode(x.cuda())
            want_q = m.decode(want)
        assert torch.equal(codes_h, want.cpu())
        assert torch.equal(quant_h, want_q.transpose(1, 2).cpu())
