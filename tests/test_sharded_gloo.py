"""world_size-2 gloo tests (CPU) of the frame-sharding host logic in audiocraft_b200/sharded.py.

The product kernels need CUDA, so the per-rank compute is played by a stand-in quantizer backed by
the CPU oracle (tests may use the oracle); what is under test is the partitioning, the codebook
broadcast and the all-gather layout."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
from audiocraft_b200.sharded import ShardedRVQ, shard_axis, shard_bounds, shard_sizes
from tests.golden import gen


class OracleQuantizer(torch.nn.Module):
    """BaseQuantizer-shaped stand-in: buffers like the real module, arithmetic by the CPU oracle."""

    def __init__(self, embed, bins):
        super().__init__()
        self.bins = bins
        for i, e in enumerate(embed):
            self.register_buffer(f"embed{i}", torch.from_numpy(e.copy()))

    def _embed(self):
        return [b.numpy() for b in self.buffers()]

    def encode(self, x):
        return torch.from_numpy(np.ascontiguousarray(oracle.encode(x.numpy(), self._embed())))

    def decode(self, codes):
        return torch.from_numpy(np.ascontiguousarray(oracle.decode(codes.numpy(), self._embed())))


def test_shard_math():
    assert shard_sizes(10, 4) == [3, 3, 2, 2]
    assert [shard_bounds(10, 4, r) for r in range(4)] == [(0, 3), (3, 6), (6, 8), (8, 10)]
    assert shard_sizes(3, 8) == [1, 1, 1, 0, 0, 0, 0, 0]
    assert shard_axis(64, 1500, 8) == 0 and shard_axis(1, 1500, 8) == 2
    for n in (0, 1, 7, 96000, 1 << 24):
        for w in (1, 2, 4, 8):
            assert sum(shard_sizes(n, w)) == n
            assert shard_bounds(n, w, w - 1)[1] == n


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, shape, wire_int16, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        B, T, K, C, D = shape
        oracle.set_threads(2)
        embed = gen.codebooks(50, K, C, D)
        if rank != 0:  # rank 0's codebooks must win
            embed = [np.zeros_like(e) for e in embed]
        q = OracleQuantizer(embed, C)
        sh = ShardedRVQ(q, wire_int16=wire_int16)
        x = torch.from_numpy(gen.normalish(51, (B, D, T)))
        codes = sh.encode(x)
        assert codes.shape == (B, K, T) and codes.dtype == torch.int64
        # already-local entry point gives the same
        codes2 = sh.encode(sh.local_slice(x), global_shape=(B, D, T), already_local=True)
        assert torch.equal(codes, codes2)
        dec_local = sh.decode(codes)
        out[rank] = (codes.numpy().copy(), dec_local.numpy().copy(), sh.local_slice(x).shape)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("shape,wire_int16", [((5, 40, 3, 64, 32), True), ((1, 75, 2, 64, 32), False),
                                              ((4, 16, 2, 40000, 16), None)])
def test_world2_gloo_encode_gather_matches_single_process(shape, wire_int16):
    B, T, K, C, D = shape
    if C > 1000:  # exercises the int64 wire format chosen automatically for bins > 32767
        C = 40000
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    port = _free_port()
    mp.spawn(_worker, args=(world, port, (B, T, K, C, D), wire_int16, out), nprocs=world, join=True)
    embed = gen.codebooks(50, K, C, D)
    x = gen.normalish(51, (B, D, T))
    want = oracle.encode(x, embed)
    for r in range(world):
        codes, dec_local, local_shape = out[r]
        assert np.array_equal(codes, want), f"rank {r}"
    # the local decodes tile the global decode
    full = oracle.decode(want, embed)
    axis = shard_axis(B, T, world)
    got = np.concatenate([out[r][1] for r in range(world)], axis=0 if axis == 0 else 2)
    assert np.array_equal(got, full)


def _avg_worker(rank, world, port, out):
    from audiocraft_b200.quantization import ResidualVectorQuantizer
    from audiocraft_b200.sharded import average_buffers
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(10 + rank)
        m = ResidualVectorQuantizer(dimension=16, n_q=2, bins=8, kmeans_init=False)
        for b in m.buffers():
            b.add_(float(rank))
        before = {k: v.clone() for k, v in m.state_dict().items()}
        average_buffers(m)
        out[rank] = (before, {k: v.clone() for k, v in m.state_dict().items()})
    finally:
        dist.destroy_process_group()


def test_world2_gloo_average_buffers_is_the_mean_over_workers():
    """flashy.distrib.average_tensors(rvq.buffers()) (modules/conditioners.py:963) as one flat all-reduce."""
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_avg_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    b0, a0 = out[0]
    b1, a1 = out[1]
    for k in b0:
        want = (b0[k] + b1[k]) / 2
        assert torch.allclose(a0[k], want) and torch.equal(a0[k], a1[k]), k
