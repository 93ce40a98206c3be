"""Drop-in for /root/reference/code/learners.py — ActiveLearner and Committee, query half.

    ActiveLearner(estimator, query_strategy=uncertainty_sampling, ...)          code/learners.py:21-99
    Committee(learner_list, query_strategy=vote_entropy_sampling)               code/learners.py:239-416
        .vote(X)          (N, K)    float64 hard labels                          :365-381
        .vote_proba(X)    (N, K, C) float64 per-learner probabilities            :383-416
        .predict_proba(X) (N, C)    mean over learners                           :334-345
        .predict(X)       (N,)      classes_[argmax]                             :316-332

One fused GPU call scores all K learners; the float64 containers the reference builds with
`np.zeros` are kept (the per-learner values are the fp32 scores; `predict_proba` is their float64 mean, :345). BayesianOptimizer and
CommitteeRegressor (regression / training, no driver uses them) are out of scope.
"""
from __future__ import annotations

import numpy as np

from . import engine
from .base import BaseCommittee, BaseLearner
from .disagreement import vote_entropy_sampling
from .uncertainty import uncertainty_sampling


class ActiveLearner(BaseLearner):
    def __init__(self, estimator, query_strategy=uncertainty_sampling, X_training=None, y_training=None,
                 bootstrap_init=False, **fit_kwargs):
        super().__init__(estimator, query_strategy, X_training, y_training, bootstrap_init, **fit_kwargs)


class Committee(BaseCommittee):
    def __init__(self, learner_list, query_strategy=vote_entropy_sampling):
        super().__init__(learner_list, query_strategy)
        self._set_classes()

    def _set_classes(self):
        try:
            known = tuple(l.estimator.classes_ for l in self.learner_list)
        except AttributeError:
            self.classes_, self.n_classes_ = None, 0
            return
        self.classes_ = np.unique(np.concatenate(known, axis=0), axis=0)
        self.n_classes_ = len(self.classes_)

    def _add_training_data(self, X, y):  # code/learners.py:311-313
        super()._add_training_data(X, y)
        self._set_classes()

    def _widened(self, X):
        specs = self._alink_specs()
        if specs is None:
            raise TypeError("Committee needs alink_b200 learners (no host fallback)")
        per, cons = engine.score(specs, X, want_per=True, want_cons=True)
        if per.shape[2] == 1:  # code/keras_wrapper.py:263-270
            per = np.concatenate([1 - per, per], axis=2)
            cons = np.concatenate([1 - cons, cons], axis=1)
        return per, cons

    def vote_proba(self, X, **predict_proba_kwargs):
        per, _ = self._widened(X)
        return per.astype(np.float64)

    def predict_proba(self, X, **predict_proba_kwargs):
        # code/learners.py:345: np.mean over the float64 (N, K, C) container — a float64 mean of the fp32 scores,
        # not the fp32 Bagging mean of committee.py:18 (they differ by up to one fp32 ulp)
        return np.mean(self.vote_proba(X, **predict_proba_kwargs), axis=1)

    def vote(self, X, **predict_kwargs):
        per, _ = self._widened(X)
        labels = (per[:, :, 1] > per[:, :, 0]).astype(np.int64)  # argmax, first index on ties
        return self.classes_[labels].astype(np.float64)

    def predict(self, X, **predict_proba_kwargs):
        proba = self.predict_proba(X, **predict_proba_kwargs)
        return self.classes_[np.argmax(proba, axis=1)]
