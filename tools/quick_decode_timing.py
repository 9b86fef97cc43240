"""Decode-only timing on BASELINE config 4 (B=256 x T=3000, K=4): HBM-bound gather-sum."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from audiocraft_b200.quantization import ResidualVectorQuantizer
torch.manual_seed(0)
m = ResidualVectorQuantizer(dimension=128, n_q=4, bins=2048, kmeans_init=False).cuda().eval()
m.vq.check_codes = False
cs = [torch.randint(0, 2048, (256, 4, 3000), device="cuda") for _ in range(2)]
for i in range(3): m.decode(cs[i % 2])
torch.cuda.synchronize()
a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
a.record()
for i in range(20): out = m.decode(cs[i % 2])
b.record(); torch.cuda.synchronize()
us = a.elapsed_time(b) * 1e3 / 20
nbytes = 256 * 3000 * (8 * 4 + 4 * 128)
print(f"decode c4: {us:.1f} us  {nbytes / us / 1e3:.0f} GB/s  ({nbytes / us / 1e3 / 6551:.3f} of 6551 GB/s)")
