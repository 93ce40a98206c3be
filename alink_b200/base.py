"""Drop-in for /root/reference/code/base.py — BaseLearner / BaseCommittee, query half.

`query(*args, **kw)` dispatches to `query_strategy(self, *args, **kw)` exactly as code/base.py:179-195
and :295-311 do. `fit` / `teach` / `score` (code/base.py:60-160, 197-214; `ActiveLearner.teach`,
code/learners.py:81-100) keep the reference's bookkeeping and hand the batch to `estimator.fit`, i.e. the CUDA
finetune step. One deliberate difference: the reference stacks the known data with modAL's
`data_vstack`, which for the pair inputs `[X_left, X_right]` concatenates the two SIDES along axis 0 (or raises
when batch sizes differ); here each side is concatenated with itself. `BaseCommittee.teach` / `rebag` go learner
by learner as code/base.py:313-341 do.
"""
from __future__ import annotations

import numpy as np

from . import engine


def _is_pair(X):
    return isinstance(X, (list, tuple)) and len(X) == 2


def _n_rows(X):
    return len(X[0]) if _is_pair(X) else len(X)


def _take(X, idx):
    return [np.asarray(X[0])[idx], np.asarray(X[1])[idx]] if _is_pair(X) else np.asarray(X)[idx]


def _vstack(A, B):
    if _is_pair(A) != _is_pair(B):
        raise ValueError('the dimensions of the new training data and label must'
                         'agree with the training data and labels provided so far')  # code/base.py:83-84
    try:
        if _is_pair(A):
            return [np.concatenate((A[0], B[0])), np.concatenate((A[1], B[1]))]
        return np.concatenate((A, B))
    except ValueError:
        raise ValueError('the dimensions of the new training data and label must'
                         'agree with the training data and labels provided so far') from None


class BaseLearner:
    def __init__(self, estimator, query_strategy, X_training=None, y_training=None, bootstrap_init=False,
                 **fit_kwargs):
        assert callable(query_strategy), "query_strategy must be callable"  # code/base.py:52
        self.estimator = estimator
        self.query_strategy = query_strategy
        self.X_training = X_training
        self.y_training = y_training
        if X_training is not None:  # code/base.py:59-60
            self._fit_to_known(bootstrap=bootstrap_init, **fit_kwargs)

    def _alink_specs(self):
        return engine.specs_of(self.estimator)

    def _check_X_y(self, X, y):
        if _n_rows(X) != len(y):  # sklearn check_X_y(X[0], y), code/base.py:74
            raise ValueError("Found input variables with inconsistent numbers of samples: [%d, %d]" % (_n_rows(X), len(y)))

    def _add_training_data(self, X, y):
        self._check_X_y(X, y)
        if self.X_training is None:
            self.X_training, self.y_training = X, y
        else:
            self.X_training = _vstack(self.X_training, X)
            self.y_training = _vstack(np.asarray(self.y_training), np.asarray(y))

    def _fit_to_known(self, bootstrap=False, **fit_kwargs):
        X, y = self.X_training, np.asarray(self.y_training)
        if bootstrap:
            n = _n_rows(X)
            idx = np.random.choice(range(n), n, replace=True)
            X, y = _take(X, idx), y[idx]
        self.estimator.fit(X, y, **fit_kwargs)
        return self

    def _fit_on_new(self, X, y, bootstrap=False, **fit_kwargs):
        self._check_X_y(X, y)
        y = np.asarray(y)
        if bootstrap:  # the reference drops fit_kwargs on this branch (code/base.py:126-128)
            n = _n_rows(X)
            idx = np.random.choice(range(n), n, replace=True)
            self.estimator.fit(_take(X, idx), y[idx])
        else:
            self.estimator.fit(X, y, **fit_kwargs)
        return self

    def fit(self, X, y, bootstrap=False, **fit_kwargs):
        self._check_X_y(X, y)
        self.X_training, self.y_training = X, y
        return self._fit_to_known(bootstrap=bootstrap, **fit_kwargs)

    def predict(self, X, **predict_kwargs):
        return self.estimator.predict(X, **predict_kwargs)

    def predict_proba(self, X, **predict_proba_kwargs):
        return self.estimator.predict_proba(X, **predict_proba_kwargs)

    def query(self, *query_args, **query_kwargs):
        return self.query_strategy(self, *query_args, **query_kwargs)

    def score(self, X, y, **score_kwargs):
        return self.estimator.score(X, y, **score_kwargs)

    def teach(self, X, y, bootstrap=False, only_new=False, **fit_kwargs):
        """`ActiveLearner.teach` (code/learners.py:81-100)."""
        self._add_training_data(X, y)
        if not only_new:
            self._fit_to_known(bootstrap=bootstrap, **fit_kwargs)
        else:
            self._fit_on_new(X, y, bootstrap=bootstrap, **fit_kwargs)


class BaseCommittee:
    def __init__(self, learner_list, query_strategy):
        assert type(learner_list) == list, "learners must be supplied in a list"  # code/base.py:229
        self.learner_list = learner_list
        self.query_strategy = query_strategy

    def __iter__(self):
        for learner in self.learner_list:
            yield learner

    def __len__(self):
        return len(self.learner_list)

    def _alink_specs(self):
        specs = []
        for learner in self.learner_list:
            s = engine.specs_of(learner)
            if s is None:
                return None
            specs.extend(s)
        return specs

    def query(self, *query_args, **query_kwargs):
        return self.query_strategy(self, *query_args, **query_kwargs)

    def _add_training_data(self, X, y):
        for learner in self.learner_list:  # code/base.py:262-275
            learner._add_training_data(X, y)

    def _fit_to_known(self, bootstrap=False, **fit_kwargs):
        for learner in self.learner_list:
            learner._fit_to_known(bootstrap=bootstrap, **fit_kwargs)

    def _fit_on_new(self, X, y, bootstrap=False, **fit_kwargs):
        for learner in self.learner_list:
            learner._fit_on_new(X, y, bootstrap=bootstrap, **fit_kwargs)

    def fit(self, X, y, **fit_kwargs):
        for learner in self.learner_list:
            learner.fit(X, y, **fit_kwargs)
        return self

    def rebag(self, **fit_kwargs):
        self._fit_to_known(bootstrap=True, **fit_kwargs)  # code/base.py:313-324

    def teach(self, X, y, bootstrap=False, only_new=False, **fit_kwargs):
        self._add_training_data(X, y)  # code/base.py:326-341
        if not only_new:
            self._fit_to_known(bootstrap=bootstrap, **fit_kwargs)
        else:
            self._fit_on_new(X, y, bootstrap=bootstrap, **fit_kwargs)
