"""Event timeline of CTA 0 of the tensor-core encode kernel.

    python tools/tc_timeline.py build      # here (no GPU): nvcc -DRVQ_TIMELINE -> tools/micro/_build/librvq_timeline.so
    python tools/tc_timeline.py run [B]    # on the GPU box: writes gpurun_out/tc_timeline.npz and prints a digest

The timeline build is the timing build of the kernel (production code path plus clock reads) with (tag, clock)
pairs of lane 0 of every warp of CTA 0 written to the buffer passed as `acc_out`.
"""
import ctypes
import glob
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "tools", "micro", "_build", "librvq_timeline.so")


def build():
    from audiocraft_b200 import build as b
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    cmd = [b._nvcc()] + b.NVCC_FLAGS + ["-DRVQ_TIMELINE"] + sys.argv[2:] + ["-shared", "-o", OUT] + b.sources() + ["-lcudart", "-lcuda"]
    subprocess.run(cmd, check=True)
    print(OUT)


def run(B=64, T=1500):
    import numpy as np
    import torch
    from audiocraft_b200 import _lib
    from audiocraft_b200.quantization import ResidualVectorQuantizer
    torch.manual_seed(0)
    m = ResidualVectorQuantizer(dimension=128, n_q=4, bins=2048, kmeans_init=False).cuda().eval()
    x = torch.randn(B, 128, T, device="cuda")
    vq = m.vq
    Ktot, C, D = vq._shape()
    pack = vq._pack.get(vq._embeds())
    lib = ctypes.CDLL(OUT)
    for name, (restype, argtypes) in _lib.SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = restype, argtypes
    K = 4
    codes = torch.empty((K, B, T), dtype=torch.int64, device="cuda")
    info = torch.zeros((128 * 8 + 192,), device="cuda")
    tl = torch.zeros((16 * 4096 * 2 + 64,), dtype=torch.int64, device="cuda")
    ws_bytes = lib.rvq_encode_workspace_bytes(B, D, T, K, C, _lib.ALGO_TC)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
    for _ in range(3):
        tl.zero_()
        rc = lib.rvq_debug_tc_scores(x.data_ptr(), _lib.strides3(x.stride()), B, D, T, pack.data_ptr(), Ktot, C, K, 0,
                                     codes.data_ptr(), tl.data_ptr(), info.data_ptr(), ws.data_ptr(), ws_bytes,
                                     torch.cuda.current_stream().cuda_stream)
        assert rc == 0, lib.rvq_b200_last_error()
        torch.cuda.synchronize()
    ref = m.encode(x)
    if not os.environ.get("RVQ_TL_NOCHECK"):
        assert torch.equal(ref, codes.transpose(0, 1)), "timeline build changed the codes"
    hist = tl[16 * 4096 * 2:].cpu().numpy()
    print('candidates per frame (rows: stage; columns: 0,1,2,3,4,more):')
    print(hist[:32].reshape(4, 8)[:, :6])
    print('remembered groups per frame (rows: stage; columns: 0-2,3-5,6-8,9-11,12-14,15+):')
    print(hist[32:].reshape(4, 8)[:, :6])
    ev = tl[:16 * 4096 * 2].view(16, 4096, 2).cpu().numpy()
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    np.savez_compressed(os.path.join(ROOT, "gpurun_out", "tc_timeline.npz"), ev=ev)
    digest(ev)


def digest(ev):
    import numpy as np
    t0 = min(int(ev[w, 0, 1]) for w in range(ev.shape[0]) if ev[w, 0, 0])
    for w in (0, 1, 4, 8):
        n = int((ev[w, :, 0] != 0).sum())
        print(f"warp {w}: {n} events, last at {int(ev[w, n - 1, 1]) - t0}")
    # MMA warp: per item wait for B, wait for accumulator, issue
    e = ev[1]
    n = int((e[:, 0] != 0).sum())
    tags, clk = e[:n, 0], e[:n, 1] - t0
    i200 = np.where(tags == 0x200)[0]
    wb, wa, iss, slots, start = [], [], [], [], []
    for i in i200:
        if i + 3 >= n:
            break
        wb.append(clk[i + 1] - clk[i]); wa.append(clk[i + 2] - clk[i + 1]); iss.append(clk[i + 3] - clk[i + 2])
        slots.append(int(tags[i + 2]) & 1); start.append(clk[i])
    wb, wa, iss = map(np.array, (wb, wa, iss))
    print(f"mma warp: {len(wb)} items; wait_b mean {wb.mean():.0f} (sum {wb.sum()}), wait_acc/a mean {wa.mean():.0f} "
          f"(sum {wa.sum()}), issue mean {iss.mean():.0f} (sum {iss.sum()}); span {clk[n - 1]}")
    # producer
    e = ev[0]
    n = int((e[:, 0] != 0).sum())
    tags, clk = e[:n, 0], e[:n, 1] - t0
    w = clk[1::2][: n // 2] - clk[0::2][: n // 2]
    print(f"producer: {n // 2} tiles, wait_b_empty mean {w.mean():.0f} sum {w.sum()}")
    # frame warp 4 (slot 0) and 8 (slot 1): per stage phases
    for wp in (4, 8):
        e = ev[wp]
        n = int((e[:, 0] != 0).sum())
        tags, clk = e[:n, 0], e[:n, 1] - t0
        st = np.where((tags & 0xFF0) == 0x300)[0]
        print(f"frame warp {wp}: stage starts (A ready) at", [int(clk[i]) for i in st])
        for si, i in enumerate(st[:8]):
            j = st[si + 1] if si + 1 < len(st) else n
            tg, ck = tags[i:j], clk[i:j]
            a310 = ck[tg == 0x310]; a320 = ck[tg == 0x320]; a33 = ck[(tg & 0xFF0) == 0x330]
            started = (tg[(tg & 0xFF0) == 0x330] & 1)
            f350 = ck[tg == 0x350]; f360 = ck[tg == 0x360]; f370 = ck[tg == 0x370]
            first_acc = a310[0] - ck[0] if len(a310) else -1
            ldw = (a320 - a310[: len(a320)]) if len(a320) else np.array([0])
            body = (a33 - a320[: len(a33)]) if len(a33) else np.array([0])
            gaps = (a310[1:] - a33[: len(a310) - 1]) if len(a310) > 1 else np.array([0])
            print(f"  stage {si}: A->first acc {first_acc}, ld wait mean {ldw.mean():.0f}, scan body mean {body.mean():.0f}, "
                  f"gap to next acc mean {gaps.mean():.0f} max {gaps.max()}, started early {int(started.sum())}/{len(started)}, "
                  f"filter {int(f350[0] - a33[-1]) if len(f350) else -1}, rescore {int(f360[0] - f350[0]) if len(f360) else -1}, "
                  f"update {int(f370[0] - f360[0]) if len(f370) else -1}")


if __name__ == "__main__":
    if sys.argv[1] == "build":
        build()
    elif sys.argv[1] == "run":
        run(int(sys.argv[2]) if len(sys.argv) > 2 else 64)
    else:
        import numpy as np
        digest(np.load(sys.argv[2])["ev"])
