"""A/B builds of librvq_b200.so with extra -D flags, checked and timed side by side on one box.

    python tools/variant.py build NAME [-DFLAG ...]    # here (no GPU): tools/micro/_build/librvq_NAME.so
    python tools/variant.py run NAME [NAME ...]         # on the GPU box: one subprocess per variant (own timeout)
    python tools/variant.py one NAME                    # what each subprocess runs
    python tools/variant.py pytest NAME [pytest args]   # the GPU tests with the variant in place of the product library

`one` loads the variant with ctypes (the product library is not touched), encodes small shapes first (a variant
whose accumulators are wrong falls to the slow exact path: a call that takes more than 0.2 s ends the run),
compares the codes with the exact CUDA-core scan of the product library, and times BASELINE config 2 with CUDA
events (rotating inputs larger than L2).  NAME "prod" stands for audiocraft_b200/lib/librvq_b200.so.
"""
import ctypes
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
BUILD = os.path.join(ROOT, "tools", "micro", "_build")


def lib_path(name):
    if name == "prod":
        from audiocraft_b200 import _lib
        return _lib.LIB_PATH
    return os.path.join(BUILD, f"librvq_{name}.so")


def build(name, flags):
    from audiocraft_b200 import build as b
    os.makedirs(BUILD, exist_ok=True)
    out = lib_path(name)
    cmd = [b._nvcc()] + b.NVCC_FLAGS + flags + ["-shared", "-o", out] + b.sources() + ["-lcudart", "-lcuda"]
    subprocess.run(cmd, check=True)
    print(out)


def load(name):
    from audiocraft_b200 import _lib
    lib = ctypes.CDLL(lib_path(name))
    for fn_name, (restype, argtypes) in _lib.SIGNATURES.items():
        fn = getattr(lib, fn_name)
        fn.restype, fn.argtypes = restype, argtypes
    return lib


def encode_with(lib, m, x, K, algo):
    import torch
    from audiocraft_b200 import _lib
    vq = m.vq
    Ktot, C, D = vq._shape()
    pack = vq._pack.get(vq._embeds())
    B, _, T = x.shape
    codes = torch.empty((K, B, T), dtype=torch.int64, device=x.device)
    ws_bytes = lib.rvq_encode_workspace_bytes(B, D, T, K, C, algo)
    assert ws_bytes >= 0, lib.rvq_b200_last_error()
    ws = torch.empty(max(ws_bytes, 16), dtype=torch.uint8, device=x.device)
    st = torch.cuda.current_stream().cuda_stream

    def call():
        rc = lib.rvq_encode_f32(x.data_ptr(), _lib.strides3(x.stride()), B, D, T, pack.data_ptr(), Ktot, C, K,
                                codes.data_ptr(), None, ws.data_ptr(), ws_bytes, algo, st)
        assert rc == 0, lib.rvq_b200_last_error()
    return call, codes


def one(name):
    import torch
    from audiocraft_b200 import _lib
    from audiocraft_b200.quantization import ResidualVectorQuantizer
    prod = load("prod")
    lib = load(name)
    shapes = [(4, 2048, 1, 128), (4, 2048, 3, 701), (8, 1024, 1, 129), (4, 2048, 151, 256), (4, 2048, 64, 1500)]
    ok = True
    m = None
    for K, C, B, T in shapes:
        torch.manual_seed(K * 1000 + C)
        m = ResidualVectorQuantizer(dimension=128, n_q=K, bins=C, kmeans_init=False).cuda().eval()
        x = torch.randn(B, 128, T, device="cuda", generator=torch.Generator(device="cuda").manual_seed(B * T))
        call, want = encode_with(prod, m, x, K, _lib.ALGO_EXACT)
        call()
        torch.cuda.synchronize()
        call, codes = encode_with(lib, m, x, K, _lib.ALGO_TC)
        codes.fill_(-1)
        t0 = time.perf_counter()
        call()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        bad = int((codes != want).sum())
        print(f"[{name}] shape K={K} C={C} B={B} T={T}: {bad} of {codes.numel()} codes differ, first call {dt * 1e3:.2f} ms",
              flush=True)
        ok = ok and bad == 0
        if dt > 0.2:
            print(f"[{name}] far too slow (the slow exact path everywhere?): stopping", flush=True)
            return 1
    xs = [torch.randn(64, 128, 1500, device="cuda") for _ in range(4)]
    calls = [encode_with(lib, m, xx, 4, _lib.ALGO_TC)[0] for xx in xs]
    for i in range(8):
        calls[i % 4]()
    torch.cuda.synchronize()
    ts = []
    for i in range(40):
        a = torch.cuda.Event(enable_timing=True)
        b = torch.cuda.Event(enable_timing=True)
        a.record()
        calls[i % 4]()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    a = torch.cuda.Event(enable_timing=True)
    b = torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(40):
        calls[i % 4]()
    b.record()
    torch.cuda.synchronize()
    if os.environ.get("RVQ_ROUNDS"):
        # kernel time by round structure: 148 single-slot tiles, one and two full pair rounds, config 2's 750 tiles
        for label, B, T in [("148 tiles (one slot per CTA)", 148, 128), ("296 tiles (one pair round)", 296, 128),
                            ("444 tiles (pair + single)", 444, 128), ("592 tiles (two pair rounds)", 592, 128)]:
            x = torch.randn(B, 128, T, device="cuda")
            call, _ = encode_with(lib, m, x, 4, _lib.ALGO_TC)
            for _ in range(3):
                call()
            torch.cuda.synchronize()
            tt = []
            for _ in range(15):
                e0 = torch.cuda.Event(enable_timing=True)
                e1 = torch.cuda.Event(enable_timing=True)
                e0.record()
                call()
                e1.record()
                torch.cuda.synchronize()
                tt.append(e0.elapsed_time(e1))
            tt.sort()
            print(f"[{name}]   {label}: med {tt[len(tt) // 2]:.4f} ms", flush=True)
    print(f"[{name}] parity {'ok' if ok else 'FAILED'}  c2 encode: min {ts[0]:.4f} med {ts[len(ts) // 2]:.4f} "
          f"max {ts[-1]:.4f} ms; back to back {a.elapsed_time(b) / 40:.4f} ms", flush=True)
    return 0 if ok else 2


def run(names, timeout=90):
    for name in names:
        try:
            res = subprocess.run([sys.executable, os.path.abspath(__file__), "one", name], timeout=timeout)
            if res.returncode:
                print(f"[{name}] exit code {res.returncode}", flush=True)
        except subprocess.TimeoutExpired:
            print(f"[{name}] TIMEOUT after {timeout} s", flush=True)


def run_pytest(name, args):
    from audiocraft_b200 import _lib
    _lib.LIB_PATH = lib_path(name)
    import pytest
    return pytest.main(args)


if __name__ == "__main__":
    mode = sys.argv[1]
    if mode == "build":
        build(sys.argv[2], sys.argv[3:])
    elif mode == "one":
        sys.exit(one(sys.argv[2]))
    elif mode == "pytest":
        sys.exit(run_pytest(sys.argv[2], sys.argv[3:]))
    else:
        run(sys.argv[2:])
