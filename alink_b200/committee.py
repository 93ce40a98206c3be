"""Drop-in for /root/reference/code/committee.py — `Bagging(models, attacks).predict(X)` (:6-20).

`predict` = mean over the K learners of `model.predict(X)`, computed by ONE call into the CUDA path
(all K heads loop over a resident feature tile; the mean is the fused K-learner reduction) when every
model is an alink_b200 SiameseNetwork. Foreign models fall back to nothing: they raise.
`attackModel` / `resize` (image-space augmentation, :22-37) are out of scope.
"""
from __future__ import annotations

from . import engine


class Bagging:
    def __init__(self, models, attacks):
        self.models = models
        self.attacks = list(attacks)

    def _alink_specs(self):
        specs = []
        for m in self.models:
            s = engine.specs_of(m)
            if s is None:
                raise TypeError("Bagging needs alink_b200 SiameseNetwork models (no host fallback)")
            specs.extend(s)
        return specs

    def predict(self, predict_on):
        _, cons = engine.score(self._alink_specs(), predict_on, want_per=False, want_cons=True)
        return cons

    def vote_proba(self, predict_on):
        per, _ = engine.score(self._alink_specs(), predict_on, want_per=True, want_cons=False)
        return per

    def attackModel(self, *a, **k):
        raise NotImplementedError("image-space noise/attack is upstream of the pool (SURVEY.md §2 rows 10-11)")
