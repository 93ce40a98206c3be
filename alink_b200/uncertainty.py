"""Drop-in for /root/reference/code/uncertainty.py — measures and samplers, same names and signatures.

    classifier_uncertainty / classifier_margin / classifier_entropy (classifier, X, **kw)   :63-130
    uncertainty_sampling / margin_sampling / entropy_sampling
        (classifier, X, n_instances=1, random_tie_break=False, **kw) -> (query_idx, [X[0][idx], X[0][idx]])   :133-217

Every number is produced on the GPU through the C ABI:
  * if `classifier` wraps alink_b200 heads, the sampler is ONE fused call
    (score -> K-learner reduce -> measure -> top-n);
  * otherwise `classifier.predict_proba(X)` supplies the probabilities and the measure + top-n run
    on the GPU (`alink_utility_from_proba`, `alink_topk_values`).
Selection order is (utility desc, index asc) — modAL's multi_argmax (:155) with the north_star tie
rule; `random_tie_break=True` reproduces modAL's shuffled_argmax (:157) with numpy's global RNG.
The reference's quirk of returning the LEFT inputs twice (:159,187,217) is kept.
"""
from __future__ import annotations

import numpy as np
from sklearn.exceptions import NotFittedError

from . import _ffi, engine


def _n_rows(X):
    if isinstance(X, engine.PairPool):
        return len(X)
    return len(X[0]) if isinstance(X, (list, tuple)) else X.shape[0]


def _measure(classifier, X, measure, not_fitted_value, **predict_proba_kwargs):
    specs = engine.specs_of(classifier)
    ctx = X.ctx if isinstance(X, engine.PairPool) else _ffi.default_context()
    if specs is not None:
        engine.score(specs, X, want_per=False, want_cons=False, ctx=ctx)
        return ctx.utility(measure, _n_rows(X))
    try:
        proba = classifier.predict_proba(X, **predict_proba_kwargs)
    except NotFittedError:
        return np.full((_n_rows(X),), not_fitted_value, dtype=np.float64)
    proba = np.asarray(proba, dtype=np.float32)
    if proba.ndim == 2:
        proba = proba[:, None, :]
    return ctx.utility_from_proba(proba, measure)


def classifier_uncertainty(classifier, X, **predict_proba_kwargs):
    """1 - max_c p (code/uncertainty.py:63-83)."""
    return _measure(classifier, X, "uncertainty", 1.0, **predict_proba_kwargs)


def classifier_margin(classifier, X, **predict_proba_kwargs):
    """p(1) - p(2) (code/uncertainty.py:86-110)."""
    return -_measure(classifier, X, "neg_margin", 0.0, **predict_proba_kwargs)


def classifier_entropy(classifier, X, **predict_proba_kwargs):
    """-sum p ln p (code/uncertainty.py:113-130)."""
    return _measure(classifier, X, "entropy", 0.0, **predict_proba_kwargs)


def _selected_rows(X, query_idx):
    if isinstance(X, engine.PairPool):
        rows = X.left_rows(query_idx)
        return [rows, rows]
    return [X[0][query_idx], X[0][query_idx]]  # (sic) left twice, code/uncertainty.py:159


def shuffled_argmax(values, n_instances=1):
    """modAL.utils.selection.shuffled_argmax [external, SURVEY App. B]: random tie-break."""
    values = np.asarray(values)
    perm = np.random.permutation(len(values))
    order = np.argsort(values[perm], kind="mergesort")[len(values) - n_instances:]
    return perm[order]


def _sample(classifier, X, measure, n_instances, random_tie_break, not_fitted_value, **kw):
    specs = engine.specs_of(classifier)
    if random_tie_break:
        util = _measure(classifier, X, measure, not_fitted_value, **kw)
        query_idx = shuffled_argmax(util, n_instances=n_instances)
    elif specs is not None:
        query_idx, _, _ = engine.fused_query(specs, X, measure, n_instances)
    else:
        util = _measure(classifier, X, measure, not_fitted_value, **kw)
        ctx = X.ctx if isinstance(X, engine.PairPool) else _ffi.default_context()
        query_idx, _, _ = ctx.topk_values(np.asarray(util, dtype=np.float32), n_instances)
    return query_idx, _selected_rows(X, query_idx)


def uncertainty_sampling(classifier, X, n_instances=1, random_tie_break=False, **uncertainty_measure_kwargs):
    return _sample(classifier, X, "uncertainty", n_instances, random_tie_break, 1.0, **uncertainty_measure_kwargs)


def margin_sampling(classifier, X, n_instances=1, random_tie_break=False, **uncertainty_measure_kwargs):
    return _sample(classifier, X, "neg_margin", n_instances, random_tie_break, 0.0, **uncertainty_measure_kwargs)


def entropy_sampling(classifier, X, n_instances=1, random_tie_break=False, **uncertainty_measure_kwargs):
    return _sample(classifier, X, "entropy", n_instances, random_tie_break, 0.0, **uncertainty_measure_kwargs)
