"""Encode with the quantized output (the eval forward) on BASELINE config 2 through the C ABI: variants against the
product library, codes and quantized compared bit for bit.

    python tools/quick_forward_timing.py [VARIANT ...]
"""
import ctypes
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from audiocraft_b200 import _lib
from audiocraft_b200.quantization import ResidualVectorQuantizer

ref = {}
for shape in [(64, 1500, 4, 2048), (3, 701, 8, 1024), (5, 129, 3, 2048)]:
    B, T, K, C = shape
    D = 128
    torch.manual_seed(K)
    m = ResidualVectorQuantizer(dimension=D, n_q=K, bins=C, kmeans_init=False).cuda().eval()
    vq = m.vq
    pack = vq._pack.get(vq._embeds())
    xs = [torch.randn(B, D, T, device="cuda") for _ in range(4)]
    for name in ["prod"] + sys.argv[1:]:
        path = _lib.LIB_PATH if name == "prod" else os.path.join(os.path.dirname(os.path.abspath(__file__)), "micro", "_build", f"librvq_{name}.so")
        lib = ctypes.CDLL(path)
        for fn_name in ("rvq_encode_f32", "rvq_encode_workspace_bytes"):
            fn = getattr(lib, fn_name)
            fn.restype, fn.argtypes = _lib.SIGNATURES[fn_name]
        codes = torch.empty((K, B, T), dtype=torch.int64, device="cuda")
        q = torch.empty((B, T, D), device="cuda")
        ws_bytes = lib.rvq_encode_workspace_bytes(B, D, T, K, C, _lib.ALGO_TC)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
        st = torch.cuda.current_stream().cuda_stream

        def call(x, with_q=True):
            rc = lib.rvq_encode_f32(x.data_ptr(), _lib.strides3(x.stride()), B, D, T, pack.data_ptr(), K, C, K,
                                    codes.data_ptr(), q.data_ptr() if with_q else None, ws.data_ptr(), ws_bytes, _lib.ALGO_TC, st)
            assert rc == 0
        q.fill_(float("nan"))
        call(xs[0])
        torch.cuda.synchronize()
        if name == "prod":
            ref[shape] = (codes.clone(), q.clone())
        same = torch.equal(codes, ref[shape][0]) and torch.equal(q, ref[shape][1])
        res = []
        for with_q in (True, False):
            for i in range(6):
                call(xs[i % 4], with_q)
            torch.cuda.synchronize()
            a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
            a.record()
            for i in range(40):
                call(xs[i % 4], with_q)
            b.record(); torch.cuda.synchronize()
            res.append(a.elapsed_time(b) / 40)
        print(f"[{name}] B={B} T={T} K={K} C={C}: codes+quantized {res[0]:.4f} ms, codes only {res[1]:.4f} ms, same as prod: {same}", flush=True)
