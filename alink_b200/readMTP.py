"""Drop-in for `createMiniBatch(X_plain)` of /root/reference/code/readMTP.py:123-135 (all plain_i x plain_j
blocks, label 1 iff i == j) on per-identity groups of FEATURE rows: returns `(PairPool, Y)` (see pairs.py)."""
from __future__ import annotations

from . import pairs


def createMiniBatch(X_plain, ctx=None):
    return pairs.createMiniBatch(X_plain, None, ctx=ctx)
