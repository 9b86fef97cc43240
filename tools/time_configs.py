"""Times the BASELINE.json configurations that are not the bench line (informational; results go to DESIGN.md)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from audiocraft_b200.quantization import ResidualVectorQuantizer


def timeit(fn, reps=10):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


for name, K, C, B, T in [("c1 1x1500", 4, 2048, 1, 1500), ("c2 64x1500", 4, 2048, 64, 1500),
                         ("c3 K=32 C=1024 64x2250", 32, 1024, 64, 2250),
                         ("c5 shard (1/8 of 16M frames) K=8", 8, 2048, 1024, 2048)]:
    torch.manual_seed(0)
    m = ResidualVectorQuantizer(dimension=128, n_q=K, bins=C, kmeans_init=False).cuda().eval()
    m.vq.check_codes = False
    x = torch.randn(B, 128, T, device="cuda")
    with torch.no_grad():
        codes = m.encode(x)
        te = timeit(lambda: m.encode(x))
        td = timeit(lambda: m.decode(codes))
        assert torch.equal(m.decode(codes), m(x, 50).x)
    N = B * T
    flops = 2.0 * 128 * C * K * N
    print(f"{name:40s} frames {N:9d}  encode {te:8.3f} ms = {N / te / 1e3:8.1f} Mframes/s = {flops / te / 1e9:7.1f} TFLOP/s"
          f" ({flops / te / 1e9 / 1661.2:.3f} of bf16 peak)   decode {td * 1e3:8.1f} us = {N * (8 * K + 512) / td / 1e6:7.1f} GB/s")
    del x, codes, m
    torch.cuda.empty_cache()
