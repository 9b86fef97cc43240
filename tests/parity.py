"""Parity helpers shared by the CPU and GPU tests.

Policy: codes must equal the reference's, except where fp32 rounding alone decides
between two codes.  Such a near-tie is accepted only if, recomputed in fp64 on the residual both
paths shared up to that stage, the two codes' squared distances differ by at most
2 ulp_fp32(||r||^2).  After an accepted near-tie the frame's later stages see different residuals
and are not compared.
"""
from __future__ import annotations

import numpy as np


def adjudicate(x, embed, got, want, n_q=None):
    """x [B,D,T]; embed list of [C,D]; got/want codes [B,K,T].
    Returns (n_exact_equal, n_near_ties, n_wrong, n_compared)."""
    got = np.asarray(got).astype(np.int64)
    want = np.asarray(want).astype(np.int64)
    assert got.shape == want.shape, (got.shape, want.shape)
    B, K, T = want.shape
    diff = got != want
    near = wrong = 0
    skipped = 0
    for b, t in zip(*np.nonzero(diff.any(axis=1))):
        k = int(np.nonzero(diff[b, :, t])[0][0])
        r = np.asarray(x[b, :, t], dtype=np.float32).copy()
        for j in range(k):
            r = r - embed[j][want[b, j, t]]
        r64 = r.astype(np.float64)
        e64 = embed[k].astype(np.float64)
        a, c = got[b, k, t], want[b, k, t]
        da = ((r64 - e64[a]) ** 2).sum()
        dc = ((r64 - e64[c]) ** 2).sum()
        xx = np.float32((r64 ** 2).sum())
        tol = 2.0 * float(np.spacing(np.abs(xx))) if np.isfinite(xx) else 0.0
        if np.isfinite(da) and np.isfinite(dc) and abs(da - dc) <= tol:
            near += 1
        else:
            wrong += 1
        skipped += K - k - 1
    n_cmp = want.size - skipped
    n_equal = int((~diff).sum())
    return n_equal, near, wrong, n_cmp
