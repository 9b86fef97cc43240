"""Small fixed workload for ncu captures: BASELINE config 2 encode x6, then config 4 decode x4 (same process).

    python tools/ncu_target.py > gpurun_out/plain.log 2>&1 &&
    ncu --set full --clock-control none --import-source on -k regex:rvq_encode_tc -s 3 -c 1 -o gpurun_out/enc python tools/ncu_target.py
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from audiocraft_b200.quantization import ResidualVectorQuantizer

torch.manual_seed(0)
m = ResidualVectorQuantizer(dimension=128, n_q=4, bins=2048, kmeans_init=False).cuda().eval()
m.vq.check_codes = False
xs = [torch.randn(64, 128, 1500, device="cuda") for _ in range(3)]
with torch.no_grad():
    for i in range(6):
        c = m.encode(xs[i % 3])
    codes = [torch.randint(0, 2048, (256, 4, 3000), device="cuda") for _ in range(2)]
    for i in range(4):
        q = m.decode(codes[i % 2])
torch.cuda.synchronize()
print("ok", c.shape, q.shape)
