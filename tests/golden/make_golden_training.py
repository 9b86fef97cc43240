"""Generate tests/golden/rvq_golden_training.npz: three TRAINING steps of the unmodified reference per config
(codes of every step, the four buffers of every layer after the last step, and after every step for the small config).

Run in the authoring container only (needs /root/reference):
    python tests/golden/make_golden_training.py
Configs avoid the reference's random draws (kmeans_init=False, threshold_ema_dead_code=0), so the steps are pure
arithmetic: nearest-code search, EMA of usage and per-code input sums, smoothed re-normalisation
(core_vq.py:195-219), residual chain with straight-through estimator (core_vq.py:363-385).
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from tests.golden import gen  # noqa: E402

CONFIGS = {
    # tensor-core shape (D = 128, C % 128 == 0): the per-layer search runs on the tcgen05 kernel
    "tc_d128_c256": dict(D=128, C=256, K=2, B=4, T=500, seed=71, every_step=False),
    # small shape: exact scan; buffers recorded after every step
    "small_d32_c64": dict(D=32, C=64, K=3, B=2, T=300, seed=72, every_step=True),
}
STEPS = 3


def inputs(cfg):
    embed = gen.codebooks(cfg["seed"], cfg["K"], cfg["C"], cfg["D"])
    xs = [gen.normalish(cfg["seed"] * 10 + s, (cfg["B"], cfg["D"], cfg["T"])) for s in range(STEPS)]
    return embed, xs


def kwargs(cfg):
    return dict(dimension=cfg["D"], n_q=cfg["K"], bins=cfg["C"], kmeans_init=False, threshold_ema_dead_code=0, decay=0.9)


def load_codebooks(m, embed):
    sd = m.state_dict()
    for i, e in enumerate(embed):
        sd[f"vq.layers.{i}._codebook.embed"] = torch.from_numpy(e.copy())
        sd[f"vq.layers.{i}._codebook.embed_avg"] = torch.from_numpy(e.copy())
    m.load_state_dict(sd, strict=True)


def main():
    from oracle import ref_shim
    q = ref_shim.load()
    torch.set_num_threads(os.cpu_count() or 1)
    out = {}
    for name, cfg in CONFIGS.items():
        embed, xs = inputs(cfg)
        m = q.ResidualVectorQuantizer(**kwargs(cfg))
        load_codebooks(m, embed)
        m.train()
        for s in range(STEPS):
            res = m(torch.from_numpy(xs[s]), 50)
            out[f"{name}/codes{s}"] = res.codes.numpy().astype(np.int16)
            out[f"{name}/penalty{s}"] = np.array(float(res.penalty.detach()), dtype=np.float64)
            out[f"{name}/x_sha{s}"] = np.array(gen.sha(np.ascontiguousarray(res.x.detach().numpy())))
            if cfg["every_step"] or s == STEPS - 1:
                for k, v in m.state_dict().items():
                    out[f"{name}/step{s}/{k}"] = v.numpy().copy()
        print(name, "done")
    path = os.path.join(HERE, "rvq_golden_training.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes; torch", torch.__version__)


if __name__ == "__main__":
    main()
