"""Pair enumeration of the reference, as index arrays.

The reference materialises every pair with four nested Python loops and `np.stack`
(/root/reference/code/readDFW.py:222-244 `createMiniBatch(X_plain, X_dig)`: all plain_i x dig_j blocks, then
all dig_i x dig_j blocks, each block row-major with the left image outermost; code/readMTP.py:123-135
`createMiniBatch(X_plain)`: all plain_i x plain_j blocks; label 1 iff i == j), and the drivers then featurise the
copies (code/ALINK.py:151-157). Here the groups are stacked once into a row table and the same pair ORDER is
emitted as two int32 index vectors, computed with vectorised numpy (no per-pair Python work), so pair n of
the pool is pair n of the reference's batch and `Y` is the reference's `(N, 1)` label column.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import numpy as np

from . import engine


def _blocks_to_indices(a_start, a_len, b_start, b_len, label) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """Concatenate row-major blocks: block t lists (a_start[t] + r, b_start[t] + c) for r < a_len[t], c < b_len[t]."""
    a_start, a_len, b_start, b_len, label = (np.asarray(v, np.int64) for v in (a_start, a_len, b_start, b_len, label))
    size = a_len * b_len
    N = int(size.sum())
    if N == 0:
        raise ValueError("need at least one array to stack")  # what np.stack([]) raises in the reference
    blk = np.repeat(np.arange(len(size)), size)
    first = np.cumsum(size) - size
    off = np.arange(N, dtype=np.int64) - first[blk]
    bl = b_len[blk]
    r = off // bl
    ia = (a_start[blk] + r).astype(np.int32)
    ib = (b_start[blk] + off - r * bl).astype(np.int32)
    return ia, ib, label[blk].reshape(N, 1)


def _starts(sizes, base=0):
    sizes = np.asarray(sizes, np.int64)
    return base + np.cumsum(sizes) - sizes


def pair_indices_dfw(plain_sizes: Sequence[int], dig_sizes: Sequence[int]):
    """(ia, ib, Y) of code/readDFW.py:222-244 over the table [plain_0 .. plain_{P-1}, dig_0 .. dig_{G-1}]."""
    ps, ds = np.asarray(plain_sizes, np.int64), np.asarray(dig_sizes, np.int64)
    P, G = len(ps), len(ds)
    p0, d0 = _starts(ps), _starts(ds, int(ps.sum()))
    i, j = np.repeat(np.arange(P), G), np.tile(np.arange(G), P)  # plain_i x dig_j, i outer (readDFW.py:224-233)
    i2, j2 = np.repeat(np.arange(G), G), np.tile(np.arange(G), G)  # dig_i x dig_j (readDFW.py:234-243)
    return _blocks_to_indices(np.concatenate([p0[i], d0[i2]]), np.concatenate([ps[i], ds[i2]]),
                              np.concatenate([d0[j], d0[j2]]), np.concatenate([ds[j], ds[j2]]),
                              np.concatenate([i == j, i2 == j2]))


def pair_indices_mtp(sizes: Sequence[int]):
    """(ia, ib, Y) of code/readMTP.py:123-135 over the table [plain_0 .. plain_{P-1}]."""
    s = np.asarray(sizes, np.int64)
    P = len(s)
    st = _starts(s)
    i, j = np.repeat(np.arange(P), P), np.tile(np.arange(P), P)
    return _blocks_to_indices(st[i], s[i], st[j], s[j], i == j)


def _table(groups: List[np.ndarray]) -> np.ndarray:
    groups = [np.asarray(g) for g in groups]
    nonempty = [g for g in groups if g.shape[0]]
    if not nonempty:
        raise ValueError("need at least one array to stack")
    return np.ascontiguousarray(np.concatenate(nonempty, axis=0))


def createMiniBatch(X_plain, X_dig=None, ctx=None):
    """`readDFW.createMiniBatch(X_plain, X_dig)` / `readMTP.createMiniBatch(X_plain)` on per-identity groups of
    FEATURE rows (each group an (n_g, D) array): returns `(PairPool, Y)` where the reference returns
    `([X_left, X_right], Y)`. `materialise(pool)` gives the reference's arrays back."""
    if X_dig is None:
        ia, ib, Y = pair_indices_mtp([len(g) for g in X_plain])
        table = _table(list(X_plain))
    else:
        ia, ib, Y = pair_indices_dfw([len(g) for g in X_plain], [len(g) for g in X_dig])
        table = _table(list(X_plain) + list(X_dig))
    return engine.PairPool(table, ia, ib, ctx=ctx), Y


def materialise(table, ia, ib):
    """[X_left, X_right] as the reference stacks them (tests and small pools only: O(N*D) host copies)."""
    table = np.asarray(table)
    return [table[np.asarray(ia)], table[np.asarray(ib)]]
