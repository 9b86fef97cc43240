"""GPU tests of the integer plumbing around the tokenizer (csrc/rvq_codestream.cu), through the C ABI front end.

Integer work: every comparison is exact.  The wire path of a frame-sharded job is emulated on ONE GPU (the driver's
box has one): each "rank" encodes its shard with the int16 wire output into its part of a gathered buffer, the widening
kernel must give back the codes of the whole batch.  Stereo interleave and the delay pattern are compared with the
numpy restatements in oracle/codestream.py, which tests/test_codestream_oracle.py pins against the reference."""
import numpy as np
import pytest
import torch

from audiocraft_b200 import codes as cs
from audiocraft_b200 import kernels
from audiocraft_b200.quantization import ResidualVectorQuantizer
from audiocraft_b200.sharded import shard_bounds, shard_sizes
from oracle import codestream as ocs

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("B,T,W", [(7, 300, 2), (64, 1500, 8), (1, 1000, 3), (3, 129, 4)])
def test_wire_encode_and_widen_emulated_ranks(B, T, W):
    torch.manual_seed(0)
    m = ResidualVectorQuantizer(dimension=128, n_q=4, bins=2048, kmeans_init=False).cuda().eval()
    x = torch.randn(B, 128, T, device="cuda")
    K = 4
    axis = 0 if B >= W else 2
    n = B if axis == 0 else T
    smax = max(shard_sizes(n, W))
    shape = (K, smax, T) if axis == 0 else (K, B, smax)
    recv = torch.full((W,) + shape, -7, dtype=torch.int16, device="cuda")
    with torch.no_grad():
        whole = m.encode(x)                                            # [B, K, T]
        for r in range(W):
            lo, hi = shard_bounds(n, W, r)
            xr = x[lo:hi] if axis == 0 else x[:, :, lo:hi]
            wire = recv[r][:, :hi - lo] if axis == 0 else recv[r][:, :, :hi - lo]
            local = m.vq.encode(xr, n_q=K, wire=wire)                  # [K, b, t]
            want = whole[lo:hi] if axis == 0 else whole[:, :, lo:hi]
            assert torch.equal(local.transpose(0, 1), want)
            assert torch.equal(wire.to(torch.int64), local)
        full = kernels.widen_codes_i16(recv, W, K, B, T, axis)         # [K, B, T]
    assert torch.equal(full.transpose(0, 1), whole)
    # the exact scan writes the wire format too
    m.vq.encode_algo = "exact"
    wire = torch.zeros((K, B, T), dtype=torch.int16, device="cuda")
    with torch.no_grad():
        local = m.vq.encode(x, n_q=K, wire=wire)
    assert torch.equal(wire.to(torch.int64), local) and torch.equal(local.transpose(0, 1), whole)


def test_convert_and_bit_packing_round_trip():
    g = torch.Generator(device="cuda").manual_seed(3)
    codes = torch.randint(0, 2048, (5, 4, 1001), device="cuda", generator=g)
    view = codes.transpose(0, 1).contiguous().transpose(0, 1)          # the [B, K, T] view encode returns
    for dt in (torch.int16, torch.int32):
        small = kernels.convert_codes(view, dt)
        assert small.dtype == dt and small.is_contiguous() and torch.equal(small.to(torch.int64), codes)
        assert torch.equal(kernels.convert_codes(small, torch.int64), codes)
    for bits in (11, 10, 16, 3):
        c = codes % (1 << bits)
        packed = kernels.pack_codes(c, bits)
        assert packed.numel() == (c.numel() * bits + 7) // 8
        assert np.array_equal(packed.cpu().numpy(), ocs.pack_bits(c.cpu().numpy().reshape(-1), bits))
        assert torch.equal(kernels.unpack_codes(packed, c.shape, bits), c)


@pytest.mark.parametrize("per_timestep", [False, True])
def test_stereo_interleave_and_split(per_timestep):
    g = torch.Generator(device="cuda").manual_seed(4)
    kbt = torch.randint(0, 2048, (2, 4, 3, 257), device="cuda", generator=g)   # [c, K, B, T] blocks as encode writes them
    c0, c1 = kbt[0].transpose(0, 1), kbt[1].transpose(0, 1)                    # [B, K, T] views
    joint = cs.interleave_stereo_codes(c0, c1, per_timestep)
    want = ocs.stereo_interleave(c0.cpu().numpy(), c1.cpu().numpy(), per_timestep)
    assert np.array_equal(joint.cpu().numpy(), want)
    l, r = cs.get_left_right_codes(joint, per_timestep)
    assert torch.equal(l, c0) and torch.equal(r, c1)


@pytest.mark.parametrize("K,T,delays,empty_initial", [(4, 50, None, 0), (8, 1500, None, 0), (4, 33, [0, 0, 0, 0], 0),
                                                       (4, 20, [0, 1, 1, 3], 2), (3, 5, [0, 2, 7], 0)])
def test_delay_pattern_build_and_revert(K, T, delays, empty_initial):
    g = torch.Generator(device="cuda").manual_seed(5)
    z = torch.randint(0, 2048, (3, K, T), device="cuda", generator=g)
    prov = cs.DelayedPatternProvider(K, delays=delays, empty_initial=empty_initial)
    pat = prov.get_pattern(T)
    special = 2048
    for keep in (False, True):
        values, indexes, mask = pat.build_pattern_sequence(z, special, keep_only_valid_steps=keep)
        wv, wi, wm = ocs.delay_build(z.cpu().numpy(), prov.delays, 1 + empty_initial, special, keep)
        assert np.array_equal(values.cpu().numpy(), wv) and np.array_equal(indexes.cpu().numpy(), wi)
        assert np.array_equal(mask.cpu().numpy(), wm)
        back, bi, bm = pat.revert_pattern_sequence(values, special, keep_only_valid_steps=keep)
        rv, ri, rm = ocs.delay_revert(wv, prov.delays, 1 + empty_initial, special, T, keep)
        assert np.array_equal(back.cpu().numpy(), rv) and np.array_equal(bi.cpu().numpy(), ri)
        assert np.array_equal(bm.cpu().numpy(), rm)
        if not keep:
            assert torch.equal(back, z)          # the full layout holds every code
    # a shorter sequence than the pattern (generation in progress): later timesteps come back as the special token
    values, _, _ = pat.build_pattern_sequence(z, special)
    part = values[:, :, : values.shape[2] // 2]
    back, _, bm = pat.revert_pattern_sequence(part, special)
    rv, _, rm = ocs.delay_revert(part.cpu().numpy(), prov.delays, 1 + empty_initial, special, T, False)
    assert np.array_equal(back.cpu().numpy(), rv) and np.array_equal(bm.cpu().numpy(), rm)


@pytest.mark.parametrize("spin", [False, True])
@pytest.mark.parametrize("B,T,W", [(7, 300, 2), (16, 700, 8), (2, 1000, 3)])
def test_peer_exchange_emulated_ranks(B, T, W, spin):
    """The NVLink exchange (rvq_exchange_push / rvq_exchange_widen) with W ranks emulated on one GPU: W plain device
    buffers stand in for the peers' symmetric buffers, the ranks' kernels run one after the other on one stream (every
    flag a kernel waits for is already set when it starts).  Five steps, two slots: slots and flags are reused."""
    torch.manual_seed(0)
    K = 4
    m = ResidualVectorQuantizer(dimension=128, n_q=K, bins=2048, kmeans_init=False).cuda().eval()
    geo = kernels.ExchangeGeometry(W, K, B, T, nslots=2)
    bufs = [torch.zeros(geo.total_bytes, dtype=torch.uint8, device="cuda") for _ in range(W)]
    ptrs = [b.data_ptr() for b in bufs]
    scratch = [torch.zeros(256, dtype=torch.uint8, device="cuda") for _ in range(W)]
    status = [torch.zeros(2, dtype=torch.int64, device="cuda") for _ in range(W)]
    st = torch.cuda.current_stream().cuda_stream
    n = B if geo.axis == 0 else T
    for seq in range(1, 6):
        x = torch.randn(B, 128, T, device="cuda")
        with torch.no_grad():
            whole = m.encode(x)
            for r in range(W):
                lo, hi = shard_bounds(n, W, r)
                xr = x[lo:hi] if geo.axis == 0 else x[:, :, lo:hi]
                shard = geo.shard_view(bufs[r], seq % 2, r)
                wire = shard[:, :hi - lo] if geo.axis == 0 else shard[:, :, :hi - lo]
                m.vq.encode(xr, n_q=K, wire=wire)
            for r in range(W):
                kernels.exchange_push(geo, ptrs, r, seq, scratch[r], status[r], st, spin_wait=spin)
            for r in range(W):
                out = torch.empty((K, B, T), dtype=torch.int64, device="cuda")
                kernels.exchange_widen(geo, ptrs, r, seq, out, scratch[r], status[r], st, spin_wait=spin)
                assert torch.equal(out.transpose(0, 1), whole), f"step {seq}, rank {r}"
    for r in range(W):
        assert status[r].tolist() == [0, 0]


def test_peer_exchange_wait_times_out_instead_of_hanging():
    """A rank whose peers never push: the widening kernel gives up after about 50 ms and says whom it waited for."""
    geo = kernels.ExchangeGeometry(2, 4, 4, 64, nslots=2)
    bufs = [torch.zeros(geo.total_bytes, dtype=torch.uint8, device="cuda") for _ in range(2)]
    scratch = torch.zeros(256, dtype=torch.uint8, device="cuda")
    status = torch.zeros(2, dtype=torch.int64, device="cuda")
    out = torch.empty((4, 4, 64), dtype=torch.int64, device="cuda")
    kernels.exchange_widen(geo, [b.data_ptr() for b in bufs], 0, 1, out, scratch, status, torch.cuda.current_stream().cuda_stream,
                           spin_wait=True)   # (the stream-operation wait would wait for ever: that is its contract)
    torch.cuda.synchronize()
    assert status[0].item() == 2 and status[1].item() in (0, 1)


def test_decode_reads_delayed_sequences_and_int16_codes_in_place():
    torch.manual_seed(0)
    K, T = 4, 333
    m = ResidualVectorQuantizer(dimension=128, n_q=K, bins=2048, kmeans_init=False).cuda().eval()
    g = torch.Generator(device="cuda").manual_seed(6)
    z = torch.randint(0, 2048, (3, K, T), device="cuda", generator=g)
    pat = cs.DelayedPatternProvider(K).get_pattern(T)
    seq, _, _ = pat.build_pattern_sequence(z, 2048)
    with torch.no_grad():
        want = m.decode(z)
        got = cs.decode_pattern_sequence(m, seq, pat)
        got16 = cs.decode_compact(m, cs.to_int16(z))
    assert torch.equal(got, want) and got.stride() == want.stride()
    assert torch.equal(got16, want)
    with pytest.raises(IndexError):   # a truncated sequence still holds special tokens: out of range, as in the reference
        bad = seq.clone()
        bad[:, 3, 5] = 2048
        cs.decode_pattern_sequence(m, bad, pat)


def test_small_model_decode_from_shared_memory():
    """The debug model's codebooks (4 x 400 x 32 fp32 = 205 KB) are staged in shared memory once there are enough
    frames; the result must not depend on the path."""
    from tests.golden import gen
    from tests.test_gpu_parity import make_module, same_bits
    import oracle
    embed = gen.codebooks(61, 4, 400, 32)
    m = make_module(embed, 32, 400)
    codes = gen.randint(62, (16, 4, 8000), 400)          # 128 000 frames: staged path
    few = codes[:1, :, :100]                              # 100 frames: rows come from L2
    with torch.no_grad():
        big = m.decode(torch.from_numpy(codes).cuda())
        small = m.decode(torch.from_numpy(few.copy()).cuda())
    assert same_bits(np.ascontiguousarray(big[:2].cpu().numpy()), np.ascontiguousarray(oracle.decode(codes[:2], embed)))
    assert torch.equal(small, big[:1, :, :100])
