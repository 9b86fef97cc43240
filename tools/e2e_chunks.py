"""End-to-end (pinned host -> pinned host) throughput of HostPipeline for several chunk sizes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from audiocraft_b200.quantization import ResidualVectorQuantizer
from audiocraft_b200.host_pipeline import HostPipeline
torch.manual_seed(0)
B, D, T, K = 64, 128, 1500, 4
m = ResidualVectorQuantizer(dimension=D, n_q=K, bins=2048, kmeans_init=False).cuda().eval()
x = [torch.randn(B, D, T).pin_memory() for _ in range(2)]
c = [torch.empty((B, K, T), dtype=torch.int64).pin_memory() for _ in range(2)]
q = [torch.empty((B, T, D)).pin_memory() for _ in range(2)]
for chunk, slots in ((16, 3), (8, 3), (8, 4), (32, 3), (4, 4), (64, 2)):
    pipe = HostPipeline(m, torch.device("cuda"), chunk_items=chunk, slots=slots)
    for i in range(3): pipe.encode_decode(x[i % 2], c[i % 2], q[i % 2])
    pipe.synchronize(); torch.cuda.synchronize()
    a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(50): last = pipe.encode_decode(x[i % 2], c[i % 2], q[i % 2])
    torch.cuda.current_stream().wait_event(last); b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 50
    print(f"chunk {chunk:3d} items, {slots} staging slots: {ms:.3f} ms per batch = {B * T / ms / 1e3:.1f} Mframes/s")
