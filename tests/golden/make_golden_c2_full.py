"""Generate tests/golden/rvq_golden_c2_full.npz: the UNMODIFIED reference's codes for the whole of
BASELINE config 2 (n_q=4, bins=2048, dim=128, B=64 x T=1500 = 96 000 frames, 384 000 codes).

Run in the authoring container only (needs /root/reference):
    python tests/golden/make_golden_c2_full.py
Inputs are the portable seeded generators of tests/golden/gen.py (the bench's own inputs come from the CUDA
generator and cannot be regenerated on a CPU).  Stored: codes as int16 [64, 4, 1500], sha256 of the inputs and of
the reference's dequantized output.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from tests.golden import gen  # noqa: E402

EMBED_SEED, X_SEED, SHAPE = 7, 1234, (64, 128, 1500)


def inputs():
    return gen.normalish(X_SEED, SHAPE), gen.codebooks(EMBED_SEED, 4, 2048, 128)


def main():
    from oracle import ref_shim
    from tests.golden.make_golden import ref_module
    q = ref_shim.load()
    torch.set_num_threads(os.cpu_count() or 1)
    x, embed = inputs()
    m = ref_module(q, embed, 128, 2048)
    with torch.no_grad():
        codes = m.encode(torch.from_numpy(x))
        quant = m.decode(codes)
    out = {
        "codes": codes.numpy().astype(np.int16),
        "x_sha": np.array(gen.sha(x)),
        "embed_sha": np.array(gen.sha(np.stack(embed))),
        "quant_sha": np.array(gen.sha(np.ascontiguousarray(quant.numpy().transpose(0, 2, 1)))),
    }
    path = os.path.join(HERE, "rvq_golden_c2_full.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes; torch", torch.__version__)


if __name__ == "__main__":
    main()
