"""The exact CUDA-core scan (RVQ_ALGO_EXACT) for D = 128 / 256 / 512 through the C ABI: variants against the product
library, codes compared bit for bit.

    python tools/quick_exact_timing.py [VARIANT ...]
"""
import ctypes
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from audiocraft_b200 import _lib
from audiocraft_b200.quantization import ResidualVectorQuantizer

ref = {}
for shape in [(16, 1500, 4, 2048, 128), (16, 1500, 4, 1024, 256), (16, 750, 6, 1024, 512), (3, 333, 3, 1000, 96)]:
    B, T, K, C, D = shape
    torch.manual_seed(D)
    m = ResidualVectorQuantizer(dimension=D, n_q=K, bins=C, kmeans_init=False).cuda().eval()
    vq = m.vq
    pack = vq._pack.get(vq._embeds())
    x = torch.randn(B, D, T, device="cuda")
    for name in ["prod"] + sys.argv[1:]:
        path = _lib.LIB_PATH if name == "prod" else os.path.join(os.path.dirname(os.path.abspath(__file__)), "micro", "_build", f"librvq_{name}.so")
        lib = ctypes.CDLL(path)
        fn = lib.rvq_encode_f32
        fn.restype, fn.argtypes = _lib.SIGNATURES["rvq_encode_f32"]
        codes = torch.empty((K, B, T), dtype=torch.int64, device="cuda")
        q = torch.empty((B, T, D), device="cuda")
        st = torch.cuda.current_stream().cuda_stream

        def call():
            rc = fn(x.data_ptr(), _lib.strides3(x.stride()), B, D, T, pack.data_ptr(), K, C, K, codes.data_ptr(),
                    q.data_ptr(), None, 0, _lib.ALGO_EXACT, st)
            assert rc == 0
        call()
        torch.cuda.synchronize()
        if name == "prod":
            ref[shape] = (codes.clone(), q.clone())
        same = torch.equal(codes, ref[shape][0]) and torch.equal(q, ref[shape][1])
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record()
        for i in range(5):
            call()
        b.record(); torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 5
        print(f"[{name}] B={B} T={T} K={K} C={C} D={D}: {ms:.3f} ms  {2e-9 * B * T * K * C * D / ms:.1f} TFLOP/s  same as prod: {same}", flush=True)
