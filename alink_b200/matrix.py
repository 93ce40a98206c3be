"""matrix.py — the DFW all-pairs score matrix (BASELINE configs[4]).

Mirrors /root/reference/utilities/generateMatrixDFW.py:19-36: load the disguised-faces model, load the (n, D)
feature table from `processedData.npy`, and for every probe row i score it against every row j — the reference
materialises two (n, D) arrays per row and calls `predict` n times; here `alink_score_allpairs` enumerates
pair (i, j) arithmetically over the device-resident table, in blocks of probe rows.
"""
from __future__ import annotations

import numpy as np

from . import _ffi, engine


def score_matrix(model, features, column=0, rows_per_block=None, ctx=None):
    """scores[i, j] = model.predict([features[i], features[j]])[column] for all i, j (generateMatrixDFW.py:28-35:
    `out[0]` = column 0). model: siamese.SiameseNetwork (or anything exposing `_alink_specs`); features: (n, D)
    float32 array. Returns (n, n) float32."""
    ctx = ctx or _ffi.default_context()
    specs = engine.specs_of(model)
    if specs is None:
        raise TypeError("model is not an alink_b200 estimator")
    feats = np.ascontiguousarray(features, dtype=np.float32)
    n = feats.shape[0]
    g = engine._GroupCache.group_for(ctx, specs)
    slot = _ffi.MAX_SLOTS - 3
    ctx.set_table(slot, feats)
    K, Cc = len(specs), specs[0].C
    col = column if Cc > 1 else 0
    if rows_per_block is None:
        rows_per_block = max(1, min(n, (1 << 24) // max(n, 1)))  # ~16 M pairs per call
    out = np.empty((n, n), np.float32)
    for r0 in range(0, n, rows_per_block):
        nr = min(rows_per_block, n - r0)
        _, cons = ctx.score_allpairs(g, slot, slot, r0, nr, n, K, Cc, want_per=False, want_cons=True)
        out[r0:r0 + nr] = cons.reshape(nr, n, Cc)[:, :, col]
    return out


def generate_matrix(model, features_npy="processedData.npy", out_path=None, column=0):
    """File-level twin of generateMatrixDFW.py: `.npy` feature table in, `np.savetxt` score matrix out."""
    scores = score_matrix(model, np.load(features_npy), column)
    if out_path is not None:
        np.savetxt(out_path, scores)
    return scores
