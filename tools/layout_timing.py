"""Encode time for the two physical input layouts: [B, D, T] (the reference's; transposed on the fly) and
[B, T, D] (frame-major; read directly)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from audiocraft_b200.quantization import ResidualVectorQuantizer
torch.manual_seed(0)
m = ResidualVectorQuantizer(dimension=128, n_q=4, bins=2048, kmeans_init=False).cuda().eval()
m.vq.check_codes = False
for name, mk in (("[B, D, T]", lambda: torch.randn(64, 128, 1500, device="cuda")),
                 ("[B, T, D]", lambda: torch.randn(64, 1500, 128, device="cuda").transpose(1, 2))):
    xs = [mk() for _ in range(4)]
    for i in range(5): m.encode(xs[i % 4])
    torch.cuda.synchronize()
    a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(40): m.encode(xs[i % 4])
    b.record(); torch.cuda.synchronize()
    print(f"{name}: {a.elapsed_time(b) / 40:.4f} ms per encode")
