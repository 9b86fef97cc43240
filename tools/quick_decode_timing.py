"""Decode-only timing on BASELINE config 4 (B=256 x T=3000, K=4) through the C ABI, output preallocated.

    python tools/quick_decode_timing.py [VARIANT ...]     # variants built by tools/variant.py; none = the product library
"""
import ctypes
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from audiocraft_b200 import _lib
from audiocraft_b200.quantization import ResidualVectorQuantizer

B, K, T, C, D = (int(os.environ.get("DEC_B", 256)), 4, int(os.environ.get("DEC_T", 3000)), 2048, 128)
torch.manual_seed(0)
m = ResidualVectorQuantizer(dimension=D, n_q=K, bins=C, kmeans_init=False).cuda().eval()
vq = m.vq
pack = vq._pack.get(vq._embeds())
codes = [torch.randint(0, C, (B, K, T), device="cuda") for _ in range(2)]
out = torch.empty((B, T, D), device="cuda")
status = torch.zeros(2, dtype=torch.int64, device="cuda")
ref = None
for name in (sys.argv[1:] or ["prod"]):
    path = _lib.LIB_PATH if name == "prod" else os.path.join(os.path.dirname(os.path.abspath(__file__)), "micro", "_build", f"librvq_{name}.so")
    lib = ctypes.CDLL(path)
    fn = lib.rvq_decode_f32
    fn.restype, fn.argtypes = _lib.SIGNATURES["rvq_decode_f32"]
    st = torch.cuda.current_stream().cuda_stream

    def call(c):
        rc = fn(c.data_ptr(), 8, _lib.strides3(c.stride()), B, K, T, pack.data_ptr(), K, C, D, out.data_ptr(),
                _lib.strides3((T * D, 1, D)), status.data_ptr(), st)
        assert rc == 0
    for i in range(4):
        call(codes[i % 2])
    torch.cuda.synchronize()
    call(codes[0])
    torch.cuda.synchronize()
    if ref is None:
        ref = out.clone()
    ok = torch.equal(ref, out)
    ts = []
    for i in range(30):
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); call(codes[i % 2]); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    us = ts[len(ts) // 2]
    nbytes = B * T * (8 * K + 4 * D)
    print(f"[{name}] decode B={B} T={T}: med {us:.1f} us (min {ts[0]:.1f})  {nbytes / us / 1e3:.0f} GB/s  same output: {ok}", flush=True)
