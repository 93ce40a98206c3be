"""roc.py — verification curves on the all-pairs score matrix.

Mirrors /root/reference/utilities/ROC_precompute.py:19-63: the reference walks the strict upper triangle of the
7771 x 7771 DFW score matrix in Python, sorts each score into Genuine / Imposter by the mask code and the
protocol (`roc_case`), then loops over every threshold and every score. Here the whole computation is one
streaming pass on the GPU (csrc/roc.cuh: 8 bytes per pair, shared-memory histograms, exact integer counts).
"""
from __future__ import annotations

import numpy as np

from . import _ffi


def roc_counts(score_matrix, masked_matrix, thresholds, roc_case, row0=0, ctx=None):
    """(TP[t], FP[t], n_genuine, n_impostor) for a block of probe rows [row0, row0 + P) of the matrix.
    score_matrix / masked_matrix: numpy arrays or torch CUDA tensors (fp32 / int32, left on the device)."""
    ctx = ctx or _ffi.default_context()
    return ctx.roc_counts(score_matrix, masked_matrix, thresholds, roc_case, row0)


def roc_precompute(score_matrix, masked_matrix, thresholds, roc_case, ctx=None):
    """np.array([true_positive_rate, false_positive_rate]) — what ROC_precompute.py:63 writes with np.savetxt.
    Division as in :59-62 (`from __future__ import division`): ZeroDivisionError if a class is empty."""
    tp, fp, ng, ni = roc_counts(score_matrix, masked_matrix, thresholds, roc_case, 0, ctx)
    if ng == 0 or ni == 0:
        raise ZeroDivisionError("no genuine or no impostor pairs for this protocol")
    return np.array([tp.astype(np.float64) / ng, fp.astype(np.float64) / ni])
