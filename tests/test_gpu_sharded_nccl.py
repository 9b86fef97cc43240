"""Two ranks, two GPUs, NCCL: the frame-sharded path on the real kernels (skipped on a one-GPU box).

Every rank encodes its slice with the CUDA kernels, the codes are all-gathered (synchronously and on the side
stream) and must equal the single-GPU codes of the whole batch; rank 0's codebooks win the broadcast."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, B, T, out):
    from audiocraft_b200.quantization import ResidualVectorQuantizer
    from audiocraft_b200.sharded import ShardedRVQ
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        torch.manual_seed(100 + rank)  # different codebooks per rank before the broadcast
        m = ResidualVectorQuantizer(dimension=128, n_q=4, bins=2048, kmeans_init=False).to(dev).eval()
        sh = ShardedRVQ(m)
        g = torch.Generator().manual_seed(7)
        x = torch.randn(B, 128, T, generator=g).to(dev)  # the same global batch on every rank
        with torch.no_grad():
            whole = m.encode(x)                         # single-GPU answer (codebooks are rank 0's by now)
            local = sh.encode_local(sh.local_slice(x))                 # plain encode, narrowed by rvq_codes_convert
            sync = sh.gather_codes(local, (B, 4, T))
            asyn, ev = sh.gather_codes_async(local, (B, 4, T))
            torch.cuda.current_stream(dev).wait_event(ev)
            wired_ok = True
            for _ in range(3):                                         # int16 written by the encode kernel; buffers rotate
                lw = sh.encode_local(sh.local_slice(x), (B, 4, T))
                fw, ev = sh.gather_codes_async(lw, (B, 4, T))
                torch.cuda.current_stream(dev).wait_event(ev)
                wired_ok = wired_ok and bool(torch.equal(fw, whole)) and bool(torch.equal(lw, local))
        torch.cuda.synchronize(dev)
        out[rank] = (bool(torch.equal(sync, whole)), bool(torch.equal(asyn, whole)) and wired_ok, whole[:, :, :8].cpu())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("B,T", [(6, 700), (1, 1500)])  # batch axis split; time axis split
def test_two_gpu_sharded_encode_matches_single_gpu(B, T):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    world = 2
    with mp.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_worker, args=(world, _free_port(), B, T, out), nprocs=world, join=True)
        assert set(out.keys()) == {0, 1}
        for r in range(world):
            sync_ok, async_ok, _ = out[r]
            assert sync_ok, f"rank {r}: gathered codes differ from the single-GPU codes"
            assert async_ok, f"rank {r}: side-stream gather differs"
        assert torch.equal(out[0][2], out[1][2]), "ranks disagree: the codebook broadcast did not take"


def _worker_p2p(rank, world, port, B, T, out):
    from audiocraft_b200.quantization import ResidualVectorQuantizer
    from audiocraft_b200.sharded import ShardedRVQ
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        torch.manual_seed(0)
        m = ResidualVectorQuantizer(dimension=128, n_q=4, bins=2048, kmeans_init=False).to(dev).eval()
        ok = {}
        for mode in ("p2p", "nccl"):
            sh = ShardedRVQ(m, exchange=mode)
            good = True
            with torch.no_grad():
                for step in range(6):
                    x = torch.randn(B, 128, T, generator=torch.Generator().manual_seed(50 + step)).to(dev)
                    whole = m.encode(x)
                    local = sh.encode_local(sh.local_slice(x), (B, 4, T))
                    full, ev = sh.gather_codes_async(local, (B, 4, T))
                    torch.cuda.current_stream(dev).wait_event(ev)
                    good = good and bool(torch.equal(full, whole))
            torch.cuda.synchronize(dev)
            sh.check_exchange()
            ok[mode] = good
        out[rank] = ok
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("B,T", [(6, 700), (1, 1500)])
def test_two_gpu_peer_to_peer_exchange(B, T):
    """The NVLink exchange on symmetric memory (rvq_exchange_push / rvq_exchange_widen) against the NCCL all-gather and
    the single-GPU codes, six steps in a row (slots and flags are reused)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    with mp.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_worker_p2p, args=(2, _free_port(), B, T, out), nprocs=2, join=True)
        for r in range(2):
            assert out[r] == {"p2p": True, "nccl": True}, f"rank {r}: {out[r]}"
