"""Drop-in for /root/reference/code/committee.py — `Bagging(models, attacks).predict(X)` (:6-20).

`predict` = mean over the K learners of `model.predict(X)`, computed by ONE call into the CUDA path
(all K heads loop over a resident feature tile; the mean is the fused K-learner reduction) when every
model is an alink_b200 SiameseNetwork. Foreign models fall back to nothing: they raise.
`attackModel` (:28-37) drives alink_b200.noise objects over embedding pairs; `resize` (:22-26) is image-space and absent.
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

    def attackModel(self, pairs, target_size=None, target_labels=None):
        """code/committee.py:28-37: one noisy copy of the pairs per attack object -> [lefts, rights] (lists indexed by
        attack, the layout code/ALINK.py:170,219-224 reads). The attacks are alink_b200.noise objects over EMBEDDING
        pairs; the reference's `resize` to the feature model's input size is an image operation and has no counterpart
        (`target_size` is accepted and ignored)."""
        perturbed_l, perturbed_r = [], []
        for attack in self.attacks:
            left, right = attack.addPairNoise(pairs, target_labels)
            perturbed_l.append(left)
            perturbed_r.append(right)
        return [perturbed_l, perturbed_r]
