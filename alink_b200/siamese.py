"""Drop-in for /root/reference/code/siamese.py (softmax-2 head) — query-selection path only.

Same class name, constructor and `predict` contract as the reference
(`SiameseNetwork(shape, modelName, learningRate)`, code/siamese.py:19-35; `predict([X_left, X_right])
-> (N, 2) float32`, code/siamese.py:130-131; `maybeLoadFromMemory() -> bool`, :114-119; `save`, :121-125).
`predict` runs the CUDA head through the C ABI; there is no host fallback. Training entry points
(`trainModel`, `finetune`, `customTrainModel`, code/siamese.py:44-112) are outside the path this
package replaces and raise NotImplementedError.

Weights: Keras Dense layout, W (in, out) row-major fp32, b (out,). h5py is not available in this
image, so checkpoints are `<modelName>.npz` with keys W1,b1,W2,b2,W3,b3 (Keras `.h5` import is a
"next" item).
"""
from __future__ import annotations

import numpy as np

from . import engine


def _glorot_uniform(rng, fan_in, fan_out):
    lim = np.sqrt(6.0 / (fan_in + fan_out))  # Keras 2.1.2 Dense default kernel_initializer
    return rng.uniform(-lim, lim, size=(fan_in, fan_out)).astype(np.float32)


class SiameseNetwork:
    _C = 2
    _ACT = "softmax"
    _H = (512, 64)  # code/siamese.py:29-30

    def __init__(self, shape, modelName, learningRate=1.0, feat_mode=0, hidden=None, seed=None):
        self.learningRate = learningRate
        self.modelName = modelName
        self.shape = tuple(shape) if not np.isscalar(shape) else (int(shape),)
        D = int(self.shape[-1])
        H1, H2 = hidden or self._H
        self._spec = engine.HeadSpec(D * (2 if feat_mode else 1), H1, H2, self._C, self._ACT, feat_mode)
        rng = np.random.default_rng(seed)
        s = self._spec
        self._spec.set(_glorot_uniform(rng, s.D_in, H1), np.zeros(H1, np.float32),
                       _glorot_uniform(rng, H1, H2), np.zeros(H2, np.float32),
                       _glorot_uniform(rng, H2, self._C), np.zeros(self._C, np.float32))

    # ---- weights
    def set_weights(self, W1, b1, W2, b2, W3, b3):
        self._spec.set(W1, b1, W2, b2, W3, b3)

    def get_weights(self):
        s = self._spec
        return [s.W1, s.b1, s.W2, s.b2, s.W3, s.b3]

    def maybeLoadFromMemory(self):
        try:
            z = np.load(self.modelName + ".npz")
            self._spec.set(z["W1"], z["b1"], z["W2"], z["b2"], z["W3"], z["b3"])
            return True
        except Exception:  # the reference swallows every failure (code/siamese.py:118-119)
            return False

    def save(self, customName=None):
        s = self._spec
        np.savez((customName or self.modelName) + ".npz", W1=s.W1, b1=s.b1, W2=s.W2, b2=s.b2, W3=s.W3, b3=s.b3)

    # ---- inference (the path)
    def preprocess(self, X):
        return X

    def predict(self, X):
        per, _ = engine.score([self._spec], self.preprocess(X), want_per=True, want_cons=False)
        return per[:, 0, :]

    def _alink_specs(self):
        return [self._spec]

    # ---- out of scope
    def _no_training(self, *a, **k):
        raise NotImplementedError("training is outside the query-selection path (SURVEY.md §8f item 2)")

    trainModel = finetune = customTrainModel = testAccuracy = getDenseBarebones = _no_training


class SmallRes(SiameseNetwork):
    """Dense head of the reference's low-resolution model (128/32/2, code/siamese.py:162-167) applied to
    pre-computed conv features; the conv encoder itself is upstream of the pool (out of scope)."""
    _H = (128, 32)
