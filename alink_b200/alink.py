"""The A-LINK selector (paper rule) — the math of /root/reference/code/ALINK.py:170-201 on scores.

`disparity_select(p2, p1, ...)`: p2 (T, N) = column `col` of the disguised-faces model M2 on each of
the T noisy copies, p1 (N,) = the same column of the committee mean on the clean pool. Runs on the
GPU (csrc/disparity.cuh). `alink_select(...)` is the convenience wrapper that also does the scoring:
committee on clean features, M2 on each noisy copy (code/ALINK.py:160,170).
"""
from __future__ import annotations

import numpy as np

from . import _ffi, engine


def disparity_select(p2, p1, disparity_ratio=0.25, eps=0.05, blind_strategy=False, y=None, ctx=None):
    """Returns (queryIndices ascending, ACTIVE_COUNT increment) — code/ALINK.py:172-201."""
    ctx = ctx or _ffi.default_context()
    return ctx.disparity_select(np.asarray(p2, np.float32), np.asarray(p1, np.float32), disparity_ratio, eps,
                                blind_strategy, y)


def alink_select(bag, disguised_model, clean_X, noisy_Xs, col=1, disparity_ratio=0.25, eps=0.05,
                 blind_strategy=False, y=None):
    """bag.predict(clean) + M2.predict(noisy_t) for each t, then the disparity rule.
    col = 1 in ALINK.py (:175-176), 0 in ALINK_arc.py (:172-173) / ALINK_MTP.py (:183-184)."""
    ens = bag.predict(clean_X)
    p1 = np.ascontiguousarray(ens[:, col if ens.shape[1] > 1 else 0])
    p2 = []
    for Xt in noisy_Xs:
        dp = disguised_model.predict(Xt)
        p2.append(dp[:, col if dp.shape[1] > 1 else 0])
    idx, active = disparity_select(np.stack(p2), p1, disparity_ratio, eps, blind_strategy, y)
    return idx, active, ens
