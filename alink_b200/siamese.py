"""Drop-in for /root/reference/code/siamese.py (softmax-2 head) — query-selection path only.

Same class name, constructor and `predict` contract as the reference
(`SiameseNetwork(shape, modelName, learningRate)`, code/siamese.py:19-35; `predict([X_left, X_right])
-> (N, 2) float32`, code/siamese.py:130-131; `maybeLoadFromMemory() -> bool`, :114-119; `save`, :121-125).
`predict` runs the CUDA head through the C ABI; there is no host fallback. Training entry points
(`trainModel`, `finetune`, `customTrainModel`, code/siamese.py:44-112) are outside the path this
package replaces and raise NotImplementedError.

Weights: Keras Dense layout, W (in, out) row-major fp32, b (out,). Checkpoints are the reference's
`<modelName>.h5` Keras weight files, read and written by `keras_h5` (a self-contained HDF5-subset
reader/writer: h5py is not in this image); `<modelName>.npz` with keys W1,b1,W2,b2,W3,b3 is read as a
fallback when no `.h5` exists.
"""
from __future__ import annotations

import numpy as np

import os

from . import engine, keras_h5


# Keras 2.1.2 (the reference's pin, requirements.txt:18): Adadelta(lr=1.0, rho=0.95, epsilon=1e-8) — the switch to
# K.epsilon() = 1e-7 came with 2.1.3 [external: keras/optimizers.py at tag 2.1.2, not checkable in this image]
ADADELTA_EPS = 1e-8


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
            s = self._spec
            if os.path.exists(self.modelName + ".h5"):  # code/siamese.py:116
                s.set(*keras_h5.load_head_weights(self.modelName + ".h5", s.D_in, s.H1, s.H2, s.C))
            else:
                z = np.load(self.modelName + ".npz")
                s.set(z["W1"], z["b1"], z["W2"], z["b2"], z["W3"], z["b3"])
            return True
        except Exception:  # the reference swallows every failure (code/siamese.py:118-119)
            return False

    def save(self, customName=None):
        s = self._spec  # code/siamese.py:121-125: save_weights(<name> + ".h5")
        keras_h5.save_head_weights((customName or self.modelName) + ".h5", [s.W1, s.b1, s.W2, s.b2, s.W3, s.b3], out_act=self._ACT)

    # ---- inference (the path)
    def preprocess(self, X):
        return X

    def predict(self, X):
        per, _ = engine.score([self._spec], self.preprocess(X), want_per=True, want_cons=False)
        return per[:, 0, :]

    def _alink_specs(self):
        return [self._spec]

    # ---- the finetune ("teach") step on the selected pairs
    def fit(self, x, y, batch_size=32, epochs=1, verbose=1, validation_split=0.0, shuffle=True, **ignored):
        """Keras `Model.fit` on the head, as `KerasClassifier.fit` / `ActiveLearner.teach` reach it
        (code/keras_wrapper.py:139-165, code/base.py:107-128; e.g. `learner.teach(X=..., y=..., only_new=True,
        verbose=1, epochs=2, validation_split=0.1)`, code/existing_al.py:111-117): Keras defaults (batch 32, one
        epoch, no hold-out), no callbacks. `y`: class indices, an (N,1) column or one-hot rows."""
        y = np.asarray(y)
        if y.ndim == 2 and y.shape[1] > 1:
            y = np.argmax(y, axis=1)
        return self.finetune(x, y, epochs, batch_size, verbose, validation_split, _callbacks=False)

    def finetune(self, X, Y, epochs, batch_size, verbose=1, validation_split=0.2, _callbacks=True):
        """code/siamese.py:52-58: Keras `fit(X, to_categorical(Y), batch_size, epochs, validation_split=0.2,
        callbacks=[EarlyStopping('val_loss', min_delta=0.1, patience=5), ReduceLROnPlateau('val_loss', factor=0.2,
        patience=5, min_lr=0.01)])` with Adadelta(self.learningRate, rho 0.95, epsilon 1e-8) and binary cross-entropy. [external: Keras 2.1.2
        fit semantics] the LAST `validation_split` of the samples is held out, the rest is reshuffled every epoch
        with np.random.shuffle (seeded by the driver, code/ALINK.py:19) and cut into batches of `batch_size`.
        Every pass is one `alink_train_epoch` (= the batches' `alink_train_step`s back to back) on device-resident weights;
        returns a Keras-History-like dict."""
        from . import _ffi
        XL, XR = engine._as_pair(self.preprocess(X))
        y = np.asarray(Y, dtype=np.float32).reshape(-1)
        n = len(y)
        if n != len(XL):
            raise ValueError("X and Y disagree on the number of pairs")
        ctx = _ffi.default_context()
        sa, sb = _ffi.MAX_SLOTS - 2, _ffi.MAX_SLOTS - 1
        ctx.set_table(sa, XL)
        ctx.set_table(sb, XR)
        g = engine._GroupCache.group_for(ctx, [self._spec])
        spec = self._spec
        shapes = [(spec.D_in, spec.H1), (spec.H1,), (spec.H1, spec.H2), (spec.H2,), (spec.H2, spec.C), (spec.C,)]
        if getattr(self, "_opt", None) is not None:  # Adadelta accumulators live as long as the model (Keras optimizer)
            ctx.set_train_state(g, 0, 1, self._opt[0])
            ctx.set_train_state(g, 0, 2, self._opt[1])
        split = int(n * (1.0 - validation_split)) if validation_split else n
        tr_idx, va_idx = np.arange(split, dtype=np.int32), np.arange(split, n, dtype=np.int32)
        lr = float(self.learningRate)
        hist = {"loss": [], "acc": [], "val_loss": [], "val_acc": [], "lr": []}
        es_best, es_wait = np.inf, 0
        rl_best, rl_wait = np.inf, 0

        def run(idx, apply):
            if not len(idx):
                return float("nan"), float("nan")
            idx = np.ascontiguousarray(idx, dtype=np.int32)
            # one call per pass: the batches are enqueued back to back on the device (alink_train_epoch), bit-identical
            # to a loop of alink_train_step; the sample-weighted means are formed here as Keras' fit loop forms them
            ls, accs = ctx.train_epoch(g, 0, sa, sb, idx, idx, y[idx], batch_size, lr=lr, eps=ADADELTA_EPS, apply=apply)
            tot_l = tot_a = 0.0
            for i, lo in enumerate(range(0, len(idx), batch_size)):
                nb = min(batch_size, len(idx) - lo)
                tot_l += float(ls[i]) * nb
                tot_a += float(accs[i]) * nb
            return tot_l / len(idx), tot_a / len(idx)

        order = tr_idx.copy()  # Keras' _fit_loop shuffles ONE index_array in place, epoch after epoch [external: Keras 2.1.2]
        for epoch in range(int(epochs)):
            np.random.shuffle(order)
            tl, ta = run(order, True)
            vl, va = run(va_idx, False)
            for k, v in zip(("loss", "acc", "val_loss", "val_acc", "lr"), (tl, ta, vl, va, lr)):
                hist[k].append(v)
            if verbose:
                print("Epoch %d/%d - loss: %.4f - acc: %.4f - val_loss: %.4f - val_acc: %.4f" % (epoch + 1, epochs, tl, ta, vl, va))
            if len(va_idx) and _callbacks:
                # EarlyStopping(monitor='val_loss', min_delta=0.1, patience=5)
                if vl + 0.1 < es_best:
                    es_best, es_wait = vl, 0
                else:
                    es_wait += 1
                    if es_wait >= 5:
                        break
                # ReduceLROnPlateau(monitor='val_loss', factor=0.2, patience=5, min_lr=0.01, epsilon=1e-4)
                if vl < rl_best - 1e-4:
                    rl_best, rl_wait = vl, 0
                else:
                    if rl_wait >= 5:
                        if lr > 0.01:
                            lr = max(lr * 0.2, 0.01)
                        rl_wait = 0
                    rl_wait += 1
        new = ctx.get_train_state(g, 0, 0, shapes)
        self._opt = (ctx.get_train_state(g, 0, 1, shapes), ctx.get_train_state(g, 0, 2, shapes))
        self._spec.set(*new)  # version bump: the next scoring call re-uploads heads, operand images and bounds
        self.learningRate = lr
        return hist

    def testAccuracy(self, X, Y, batch_size=512):
        """code/siamese.py:60-79: accuracy of argmax(predict) against 1*(Y[i] == Y[j]) over ALL ordered pairs
        (i, j) of X — one all-pairs scoring call instead of n^2 / batch_size predict calls."""
        from . import _ffi
        feats = np.ascontiguousarray(self.preprocess(X), dtype=np.float32)
        lab = np.asarray(Y).reshape(-1)
        n = len(feats)
        ctx = _ffi.default_context()
        g = engine._GroupCache.group_for(ctx, [self._spec])
        slot = _ffi.MAX_SLOTS - 3
        ctx.set_table(slot, feats)
        rows = max(1, min(n, (1 << 23) // max(n, 1)))
        n_correct = 0
        for r0 in range(0, n, rows):
            nr = min(rows, n - r0)
            _, cons = ctx.score_allpairs(g, slot, slot, r0, nr, n, 1, self._C, want_per=False, want_cons=True)
            p = cons.reshape(nr, n, self._C)
            pred = np.argmax(p, axis=2) if self._C == 2 else (p[:, :, 0] > 0.5).astype(np.int64)
            n_correct += int(np.sum(pred == (lab[r0:r0 + nr, None] == lab[None, :]).astype(np.int64)))
        return n_correct / float(n * n)

    # ---- out of scope: generator-driven pre-training (data loaders, class weights: code/siamese.py:44-50, 81-112)
    def _no_training(self, *a, **k):
        raise NotImplementedError("generator-driven pre-training is outside the query-selection path; "
                                  "finetune() / testAccuracy() are implemented (SURVEY.md §8f item 2)")

    trainModel = customTrainModel = getDenseBarebones = _no_training


class SmallRes(SiameseNetwork):
    """Dense head of the reference's low-resolution model (128/32/2, code/siamese.py:162-167) applied to
    pre-computed conv features; the conv encoder itself is upstream of the pool (out of scope)."""
    _H = (128, 32)
