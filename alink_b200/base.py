"""Drop-in for /root/reference/code/base.py — BaseLearner / BaseCommittee, query half.

`query(*args, **kw)` dispatches to `query_strategy(self, *args, **kw)` exactly as code/base.py:179-195
and :295-311 do. `teach` / `fit` (code/base.py:107-128, 326-341) belong to the training step that
follows the path and raise NotImplementedError.
"""
from __future__ import annotations

from . import engine


class BaseLearner:
    def __init__(self, estimator, query_strategy, X_training=None, y_training=None, bootstrap_init=False,
                 **fit_kwargs):
        assert callable(query_strategy), "query_strategy must be callable"  # code/base.py:52
        self.estimator = estimator
        self.query_strategy = query_strategy
        self.X_training = X_training
        self.y_training = y_training
        if X_training is not None:
            raise NotImplementedError("initial fitting is outside the query-selection path")

    def _alink_specs(self):
        return engine.specs_of(self.estimator)

    def predict(self, X, **predict_kwargs):
        return self.estimator.predict(X, **predict_kwargs)

    def predict_proba(self, X, **predict_proba_kwargs):
        return self.estimator.predict_proba(X, **predict_proba_kwargs)

    def query(self, *query_args, **query_kwargs):
        return self.query_strategy(self, *query_args, **query_kwargs)

    def teach(self, *a, **k):
        raise NotImplementedError("teach/fit is the step after the path (SURVEY.md §8f item 2)")

    fit = teach


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

    def teach(self, *a, **k):
        raise NotImplementedError("teach/fit is the step after the path (SURVEY.md §8f item 2)")

    fit = rebag = teach
