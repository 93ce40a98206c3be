"""Drop-in for /root/reference/code/keras_wrapper.py — the sklearn-style adapter, inference half.

`KerasClassifier(build_fn)` wraps an already built model (build_self in the ctor,
code/keras_wrapper.py:71,127-137). `predict_proba` widens a 1-column sigmoid output to
[1-p, p] (:263-270); `predict` is argmax / `> 0.5` mapped through `classes_` (:235-242).
`fit` / `score` are outside the path.
"""
from __future__ import annotations

import numpy as np

from . import engine


class KerasClassifier:
    def __init__(self, build_fn=None, **sk_params):
        self.build_fn = build_fn
        self.sk_params = sk_params
        self.model = build_fn() if callable(build_fn) else build_fn
        # a wrapped SiameseNetwork, or the reference idiom `lambda: siamese_net.siamese_net`
        self.classes_ = np.arange(2)
        self.n_classes_ = 2

    def _alink_specs(self):
        return engine.specs_of(self.model)

    def predict_proba(self, x, **kwargs):
        probs = self.model.predict(x)
        if probs.shape[1] == 1:
            probs = np.hstack([1 - probs, probs])
        return probs

    def predict(self, x, **kwargs):
        proba = self.model.predict(x)
        if proba.shape[-1] > 1:
            classes = proba.argmax(axis=-1)
        else:
            classes = (proba > 0.5).astype("int32")
        return self.classes_[classes]

    def fit(self, *a, **k):
        raise NotImplementedError("training is outside the query-selection path")
