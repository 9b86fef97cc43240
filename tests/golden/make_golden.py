"""Generate tests/golden/rvq_golden.npz by running the UNMODIFIED reference.

Run in the authoring container only (needs /root/reference):
    python tests/golden/make_golden.py
The reference's `audiocraft.quantization.ResidualVectorQuantizer` is imported through
oracle/ref_shim.py, given the seeded codebooks of tests/golden/gen.py through its own
state_dict keys (`vq.layers.{i}._codebook.embed`, loaders.py:176), put in eval mode, and asked for
encode / decode / forward.  Stored per case: the codes (int16), hashes of the inputs, and the
quantized output (full for small cases, sha256 + a strided sample for large ones).
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import ref_shim  # noqa: E402
from tests.golden import gen  # noqa: E402


def ref_module(q, embed, D, C):
    m = q.ResidualVectorQuantizer(dimension=D, n_q=len(embed), bins=C, kmeans_init=False)
    sd = m.state_dict()
    for i, e in enumerate(embed):
        sd[f"vq.layers.{i}._codebook.embed"] = torch.from_numpy(e.copy())
        sd[f"vq.layers.{i}._codebook.embed_avg"] = torch.from_numpy(e.copy())
    m.load_state_dict(sd, strict=True)
    return m.eval()


def main():
    q = ref_shim.load()
    torch.set_num_threads(os.cpu_count() or 1)
    out = {}
    for name in gen.CASES:
        case = gen.build_case(name)
        cfg = case["cfg"]
        m = ref_module(q, case["embed"], cfg["D"], cfg["C"])
        m.set_num_codebooks(case["n_q"])
        x = torch.from_numpy(case["x"])  # keeps the (possibly non-contiguous) strides
        with torch.no_grad():
            codes = m.encode(x)
            quant = m.decode(codes)
            res = m(x, 50)
        assert torch.equal(res.codes, codes), name
        same_fwd = bool(torch.equal(res.x, quant) or
                        (torch.isnan(res.x) == torch.isnan(quant)).all() and
                        torch.equal(torch.nan_to_num(res.x), torch.nan_to_num(quant)))
        assert same_fwd, name
        cn = codes.numpy()
        qn = np.ascontiguousarray(quant.numpy().transpose(0, 2, 1))  # physical [B, T, D]
        out[f"{name}/codes"] = cn.astype(np.int16)
        out[f"{name}/x_sha"] = np.array(gen.sha(case["x"]))
        out[f"{name}/embed_sha"] = np.array(gen.sha(np.stack(case["embed"])))
        out[f"{name}/quant_sha"] = np.array(gen.sha(qn))
        out[f"{name}/quant_strides"] = np.array(quant.stride(), dtype=np.int64)
        out[f"{name}/codes_strides"] = np.array(codes.stride(), dtype=np.int64)
        out[f"{name}/bandwidth"] = np.array(float(res.bandwidth), dtype=np.float64)
        out[f"{name}/penalty"] = np.array(float(res.penalty), dtype=np.float64)
        if qn.size <= 9000:
            out[f"{name}/quant"] = qn
        else:
            out[f"{name}/quant_sample"] = qn.reshape(-1)[::97].copy()
        print(f"{name:32s} codes {tuple(cn.shape)} quant {tuple(quant.shape)} strides {tuple(quant.stride())}")

    # decode-only cases (c4 shape, shrunk): random codes, int32 codes, fewer stages than layers
    embed = gen.codebooks(41, 4, 2048, 128)
    m = ref_module(q, embed, 128, 2048)
    codes = torch.from_numpy(gen.randint(42, (3, 4, 500), 2048))
    with torch.no_grad():
        d_full = m.decode(codes)
        d_i32 = m.decode(codes.to(torch.int32))
        d_k2 = m.decode(codes[:, :2])
    assert torch.equal(d_full, d_i32)
    out["decode_c4/quant_sha"] = np.array(gen.sha(np.ascontiguousarray(d_full.numpy().transpose(0, 2, 1))))
    out["decode_c4/quant_k2_sha"] = np.array(gen.sha(np.ascontiguousarray(d_k2.numpy().transpose(0, 2, 1))))
    out["decode_c4/quant_sample"] = np.ascontiguousarray(d_full.numpy().transpose(0, 2, 1)).reshape(-1)[::97].copy()
    # -0.0 rows decode to +0.0 (running sum starts from +0.0, core_vq.py:400)
    e2 = [np.full((8, 16), -0.0, dtype=np.float32)]
    m2 = ref_module(q, e2, 16, 8)
    with torch.no_grad():
        z = m2.decode(torch.zeros(1, 1, 4, dtype=torch.long))
    out["negzero/signbits"] = np.signbit(z.numpy()).astype(np.uint8)

    path = os.path.join(HERE, "rvq_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes; torch", torch.__version__)


if __name__ == "__main__":
    main()
