"""Drop-in for the pool-construction half of /root/reference/code/readDFW.py (module name kept so a driver only
changes its import): `createMiniBatch(X_plain, X_dig)` (:222-244) on per-identity groups of FEATURE rows returns
`(PairPool, Y)` in the reference's pair order (see pairs.py); `splitDisguiseData` (:212-219) is the same split.
Image readers, generators and augmentation of that file are upstream of the pool and absent."""
from __future__ import annotations

import numpy as np

from . import pairs


def createMiniBatch(X_plain, X_dig, ctx=None):
    return pairs.createMiniBatch(X_plain, X_dig, ctx=ctx)


def splitDisguiseData(X_dig, pre_ratio=0.5):
    X_dig_pre, X_dig_post = [], []
    for g in X_dig:
        g = np.asarray(g)
        cut = int(g.shape[0] * pre_ratio)
        X_dig_pre.append(g[:cut])
        X_dig_post.append(g[cut:])
    return (X_dig_pre, X_dig_post)
