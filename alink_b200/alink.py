"""The A-LINK selector (paper rule) — the math of /root/reference/code/ALINK.py:170-201 on scores.

`disparity_select(p2, p1, ...)`: p2 (T, N) = column `col` of the disguised-faces model M2 on each of
the T noisy copies, p1 (N,) = the same column of the committee mean on the clean pool. Runs on the
GPU (csrc/disparity.cuh). `alink_select(...)` is the convenience wrapper that also does the scoring:
committee on clean features, M2 on each noisy copy (code/ALINK.py:160,170).
"""
from __future__ import annotations

import numpy as np

from . import _ffi, engine


def disparity_select(p2, p1, disparity_ratio=0.25, eps=0.05, blind_strategy=False, y=None, ctx=None):
    """Returns (queryIndices ascending, ACTIVE_COUNT increment) — code/ALINK.py:172-201."""
    ctx = ctx or _ffi.default_context()
    return ctx.disparity_select(np.asarray(p2, np.float32), np.asarray(p1, np.float32), disparity_ratio, eps,
                                blind_strategy, y)


def alink_select(bag, disguised_model, clean_X, noisy_Xs, col=1, disparity_ratio=0.25, eps=0.05,
                 blind_strategy=False, y=None):
    """bag.predict(clean) + M2.predict(noisy_t) for each t, then the disparity rule.
    col = 1 in ALINK.py (:175-176), 0 in ALINK_arc.py (:172-173) / ALINK_MTP.py (:183-184)."""
    ens = bag.predict(clean_X)
    p1 = np.ascontiguousarray(ens[:, col if ens.shape[1] > 1 else 0])
    p2 = []
    for Xt in noisy_Xs:
        dp = disguised_model.predict(Xt)
        p2.append(dp[:, col if dp.shape[1] > 1 else 0])
    idx, active = disparity_select(np.stack(p2), p1, disparity_ratio, eps, blind_strategy, y)
    return idx, active, ens


def a2link_pool_select(ctx, group_m1, group_m2, slot, ia, ib, N, K1, C=2, T=16, seed=46, col=1, disparity_ratio=0.25, eps=0.05,
                       blind_strategy=False, y=None, ncoord=40, vmax=0.25, B=1000, scratch_slots=(3, 4), rule=True, pair_base=0):
    """The A2-LINK noise-augmented pool (BASELINE configs[3]), everything on the device:
    p1 = column `col` of the M1 committee's Bagging mean on the N base pairs (code/ALINK.py:160); for each of the T noise
    types the perturbed copies of both sides are GENERATED on the GPU (alink_perturb_rows: counter-based stream keyed on
    (seed, pair, side, t), the embedding-space analogue of the re-featurised noisy images of :163-168), scored by the
    one-learner model M2 (:170) and reduced to p2[t]; then (i) the reference's disparity rule (:171-201) and (ii) the
    top-B pairs by min_t |p2_t - p1|. Scores never visit the host. `ia` / `ib`: torch cuda int32 (or numpy) row indices
    into table slot `slot`; `pair_base` = global index of this shard's first pair (it keys the perturbation stream, so a sharded
    pool generates the same rows as the whole one). Returns (queryIndices, ACTIVE_COUNT, top_idx, top_util, (p1, p2) device tensors)."""
    import torch
    dev = torch.device("cuda", ctx.device)
    p1 = torch.empty((N,), dtype=torch.float32, device=dev)
    p2 = torch.empty((T, N), dtype=torch.float32, device=dev)
    ctx.score_indexed(group_m1, slot, slot, ia, ib, N, K1, C, want_per=False, want_cons=False)
    ctx.score_column(-1, col, out=p1)
    sl, sr = scratch_slots
    for t in range(T):
        ctx.perturb_rows(slot, ia, N, 0, t, seed, sl, pair_base=pair_base, ncoord=ncoord, vmax=vmax)
        ctx.perturb_rows(slot, ib, N, 1, t, seed, sr, pair_base=pair_base, ncoord=ncoord, vmax=vmax)
        ctx.score_indexed(group_m2, sl, sr, None, None, N, 1, C, want_per=False, want_cons=False)
        ctx.score_column(0, col, out=p2[t])
    idx, active = (ctx.disparity_select(p2, p1, disparity_ratio, eps, blind_strategy, y) if rule else (None, None))
    ctx.min_disparity(p2, p1)
    top_idx, top_util, _ = ctx.select_topk(min(B, N))
    return idx, active, top_idx, top_util, (p1, p2)


class ALinkState:
    """What the reference's loop carries from one iteration to the next (code/ALINK.py:135-141): the pairs selected so far
    for the disguised-faces model, the oracle-query count and the pool size seen."""

    def __init__(self):
        self.train_left, self.train_right, self.train_y = [], [], []
        self.ACTIVE_COUNT = 0
        self.UN_SIZE = 0
        self.finetunes = 0

    def pending(self):
        return int(sum(len(y) for y in self.train_y))


def roundoff(v):
    """code/helpers.py:39-46: (n, 1) hard labels of scores at 0.5."""
    return (np.asarray(v, dtype=np.float32) >= 0.5).astype(np.int64).reshape(-1, 1)


def alink_iteration(bag, disguised_model, batch_features, batch_y, state, target_size=None, col=1, disparity_ratio=0.25,
                    eps=0.05, blind_strategy=False, batch_send=64, ft_epochs=3, old_data=None, verbose=0):
    """The body of the A-LINK loop, code/ALINK.py:155-250, on a batch of EMBEDDING pairs (the reference's
    `conversionModel.process` calls of :158,167 are upstream of this package): committee prediction (:160), one noisy
    copy per noise object (`bag.attackModel`, :163-164), the disguised-faces model on each (:170), the disparity rule
    with the grey zone and the pseudo-oracle (:171-201), the per-noise partition of the selected pairs (:216-225), and —
    once `batch_send` pairs are waiting — the finetune on them plus the clean selected pairs and `old_data` (:227-250).
    `batch_y`: (N,) or (N, k) pseudo-oracle labels (column 0 is read, :195). Returns (queryIndices, ensemblePredictions)."""
    XL, XR = engine._as_pair(batch_features)
    y = np.asarray(batch_y, dtype=np.float32)
    y = y[:, 0] if y.ndim == 2 else y
    state.UN_SIZE += len(XL)
    ens = bag.predict([XL, XR])
    m1_labels = np.eye(2, dtype=np.float32)[np.argmax(ens, axis=1)]
    noisy = bag.attackModel([XL, XR], target_size, m1_labels)
    T = len(bag.attacks)
    cidx = col if ens.shape[1] > 1 else 0
    p2 = []
    for jj in range(T):
        dp = disguised_model.predict([noisy[0][jj], noisy[1][jj]])
        p2.append(dp[:, col if dp.shape[1] > 1 else 0])
    idx, active = disparity_select(np.stack(p2), np.ascontiguousarray(ens[:, cidx]), disparity_ratio, eps, blind_strategy, y)
    state.ACTIVE_COUNT += int(active)
    if len(idx) == 0:
        return idx, ens
    intermediate = ens[idx, cidx]
    mp = int(len(intermediate) / float(T))
    hard = roundoff(intermediate)
    for i in range(T):
        part = idx[i * mp:(i + 1) * mp]
        state.train_left.append(np.asarray(noisy[0][i])[part])
        state.train_right.append(np.asarray(noisy[1][i])[part])
        state.train_y.append(hard[i * mp:(i + 1) * mp])
    if state.pending() >= batch_send:
        left = state.train_left + [XL[idx]]
        right = state.train_right + [XR[idx]]
        labels = state.train_y + [hard]
        if old_data is not None:  # unperturbed pairs against overfitting on noise (:229-234)
            (ol, orr), oy = old_data
            left.append(np.asarray(ol, np.float32))
            right.append(np.asarray(orr, np.float32))
            labels.append(np.asarray(oy).reshape(-1, 1))
        disguised_model.finetune([np.concatenate(left), np.concatenate(right)], np.concatenate(labels), ft_epochs, 16, verbose)
        state.train_left, state.train_right, state.train_y = [], [], []
        state.finetunes += 1
    return idx, ens
