"""Per-role cycle counters of CTA 0 of the tensor-core encode kernel (diagnostics build)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from audiocraft_b200 import _lib
from audiocraft_b200.quantization import ResidualVectorQuantizer

B, T = int(sys.argv[1]) if len(sys.argv) > 1 else 64, 1500
if len(sys.argv) > 2:  # a variant library built by tools/variant.py
    _lib.LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "micro", "_build", f"librvq_{sys.argv[2]}.so")
torch.manual_seed(0)
m = ResidualVectorQuantizer(dimension=128, n_q=4, bins=2048, kmeans_init=False).cuda().eval()
x = torch.randn(B, 128, T, device="cuda")
lib = _lib.load()
vq = m.vq
Ktot, C, D = vq._shape()
pack = vq._pack.get(vq._embeds())
K = 4
codes = torch.empty((K, B, T), dtype=torch.int64, device="cuda")
acc = torch.zeros((128, C), device="cuda")
info = torch.zeros((128 * 8 + 192,), device="cuda")
ws_bytes = lib.rvq_encode_workspace_bytes(B, D, T, K, C, _lib.ALGO_TC)
ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
for _ in range(3):
    rc = lib.rvq_debug_tc_scores(x.data_ptr(), _lib.strides3(x.stride()), B, D, T, pack.data_ptr(), Ktot, C, K, 0,
                                 codes.data_ptr(), None, info.data_ptr(), ws.data_ptr(), ws_bytes,
                                 torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "debug")
    torch.cuda.synchronize()
t = info[1024:].view(torch.int64).view(12, 8).cpu().numpy()
a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
a.record()
lib.rvq_debug_tc_scores(x.data_ptr(), _lib.strides3(x.stride()), B, D, T, pack.data_ptr(), Ktot, C, K, 0,
                        codes.data_ptr(), None, info.data_ptr(), ws.data_ptr(), ws_bytes,
                        torch.cuda.current_stream().cuda_stream)
b.record(); torch.cuda.synchronize()
print("timing build kernel: %.3f ms" % a.elapsed_time(b))
names = {0: ["loop", "wait b_empty"], 1: ["loop", "wait b_full", "wait a_ready", "wait acc_empty", "", "", "", ""]}
print("producer  :", dict(zip(["issue", "wait_b_empty"], t[0][:2])))
print("mma       :", dict(zip(["issue", "wait_b_full", "wait_a_ready", "wait_acc_empty"], t[1][:4])))
for w in range(4, 12):
    print(f"frame w{w}  :", dict(zip(["between", "tile_load", "prep", "wait_acc", "scan", "phase1", "phase2a", "phase2b"], t[w])))
print("total cycles per warp:", t.sum(1))
