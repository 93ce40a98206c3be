"""Drop-in for /root/reference/code/keras_wrapper.py — the sklearn-style adapter, inference half.

`KerasClassifier(build_fn)` wraps an already built model (build_self in the ctor,
code/keras_wrapper.py:71,127-137). `predict_proba` widens a 1-column sigmoid output to
[1-p, p] (:263-270); `predict` is argmax / `> 0.5` mapped through `classes_` (:235-242).
`fit` (:195-218 + BaseWrapper.fit :139-165) derives `classes_` from the labels and runs the model's Keras-style
`fit` (the CUDA finetune step); `score` = mean accuracy (:272-305).
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

    def fit(self, x, y, sample_weight=None, **kwargs):
        y = np.array(y)
        if len(y.shape) == 2 and y.shape[1] > 1:
            self.classes_ = np.arange(y.shape[1])
        elif (len(y.shape) == 2 and y.shape[1] == 1) or len(y.shape) == 1:
            self.classes_ = np.unique(y)
            y = np.searchsorted(self.classes_, y)
        else:
            raise ValueError('Invalid shape for y: ' + str(y.shape))  # code/keras_wrapper.py:214
        self.n_classes_ = len(self.classes_)
        if sample_weight is not None:
            raise NotImplementedError("sample_weight is not supported by the CUDA finetune step")
        if not hasattr(self.model, "fit"):
            raise TypeError("KerasClassifier.fit needs an alink_b200 SiameseNetwork (no host fallback)")
        fit_args = dict(self.sk_params)
        fit_args.update(kwargs)
        return self.model.fit(x, y, **fit_args)

    def score(self, x, y, **kwargs):
        """Mean accuracy of `predict` (code/keras_wrapper.py:272-305 evaluates the compiled 'acc' metric)."""
        y = np.asarray(y)
        if y.ndim == 2 and y.shape[1] > 1:
            y = np.argmax(y, axis=1)
        return float(np.mean(np.asarray(self.predict(x)).reshape(-1) == y.reshape(-1)))
