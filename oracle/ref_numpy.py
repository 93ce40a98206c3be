"""ref_numpy.py — line-by-line numpy/scipy transcription of the reference's query-selection path.

TEST INFRASTRUCTURE ONLY (see oracle/alink_oracle.c header): imported by tests/, bench.py's CPU
baseline legs and tests/golden/make_golden.py; never by the product package.

It is deliberately *independent* of include/alink_math.h: it uses numpy's BLAS matmul, np.exp,
scipy.stats.entropy and np.argpartition exactly where the reference (Keras 2.1.2 / TF 1.15 /
numpy / scipy / modAL) does, so agreement between this file and the C oracle (<= 1e-6 on
probabilities) is the pin that stands in for the un-installable Keras runtime.

Each function cites the reference line(s) it transcribes (paths relative to /root/reference).
"""
from __future__ import annotations

import numpy as np
from scipy.stats import entropy as _scipy_entropy


# --------------------------------------------------------------------------- head (L2)
def head_forward(XL, XR, W1, b1, W2, b2, W3, b3, act="softmax", feat_mode=0, batch_size=1024,
                 dtype=np.float32):
    """code/siamese.py:24-33 (softmax-2) / code/siamese3.py:18-26 (sigmoid-1); batched like
    `predict(..., batch_size=1024)` (code/siamese.py:130-131)."""
    XL = np.asarray(XL, dtype=dtype)
    XR = np.asarray(XR, dtype=dtype)
    W1, b1, W2, b2, W3, b3 = (np.asarray(a, dtype=dtype) for a in (W1, b1, W2, b2, W3, b3))
    outs = []
    for s in range(0, len(XL), batch_size):
        l, r = XL[s:s + batch_size], XR[s:s + batch_size]
        x = np.abs(l - r)                                   # siamese.py:27-28  K.abs(l - r)
        if feat_mode:
            x = np.concatenate([x, l * r], axis=1)          # north_star opt-in feature (not in reference)
        h1 = np.maximum(x @ W1 + b1, 0)                     # siamese.py:29  Dense(512, relu)
        h2 = np.maximum(h1 @ W2 + b2, 0)                    # siamese.py:30  Dense(64, relu)
        z = h2 @ W3 + b3                                    # siamese.py:31  Dense(2) / siamese3.py:25 Dense(1)
        if act == "softmax":
            e = np.exp(z - z.max(axis=1, keepdims=True))    # siamese.py:32  softmax
            p = e / e.sum(axis=1, keepdims=True)
        else:
            p = 1.0 / (1.0 + np.exp(-z))                    # siamese3.py:25 sigmoid
        outs.append(p.astype(dtype))
    if not outs:
        return np.zeros((0, W3.shape[1]), dtype=dtype)
    return np.concatenate(outs, axis=0)


def bagging_predict(preds):
    """code/committee.py:13-20: np.sum(np.array(predictions), axis=0) / len(models)."""
    return np.array(np.sum(np.array(preds), axis=0) / len(preds))


def widen(p):
    """code/keras_wrapper.py:263-270: (N,1) -> hstack([1-p, p])."""
    p = np.asarray(p)
    if p.shape[1] == 1:
        p = np.hstack([1 - p, p])
    return p


def hard_labels(p):
    """code/keras_wrapper.py:235-242: argmax if >1 column else p > 0.5."""
    p = np.asarray(p)
    if p.shape[-1] > 1:
        return p.argmax(axis=-1)
    return (p > 0.5).astype("int32").reshape(-1)


def vote_proba(per_learner_probs):
    """code/learners.py:383-416: (N, K, C) float64 container of per-learner predict_proba."""
    K = len(per_learner_probs)
    P0 = widen(per_learner_probs[0])
    out = np.zeros(shape=(P0.shape[0], K, P0.shape[1]))
    for k, p in enumerate(per_learner_probs):
        out[:, k, :] = widen(p)
    return out


def committee_predict_proba(per_learner_probs):
    """code/learners.py:334-345: np.mean(vote_proba, axis=1)."""
    return np.mean(vote_proba(per_learner_probs), axis=1)


def vote(per_learner_probs):
    """code/learners.py:365-381: (N, K) float64 of hard labels."""
    K = len(per_learner_probs)
    N = len(per_learner_probs[0])
    out = np.zeros(shape=(N, K))
    for k, p in enumerate(per_learner_probs):
        out[:, k] = hard_labels(p)
    return out


# --------------------------------------------------------------------------- measures (L4)
def proba_uncertainty(proba):
    """code/uncertainty.py:26,82."""
    return 1 - np.max(proba, axis=1)


def proba_margin(proba):
    """code/uncertainty.py:40-46,104-110."""
    if proba.shape[1] == 1:
        return np.zeros(shape=len(proba))
    part = np.partition(-proba, 1, axis=1)
    return -part[:, 0] + part[:, 1]


def proba_entropy(proba):
    """code/uncertainty.py:60,130."""
    return np.transpose(_scipy_entropy(np.transpose(proba)))


def vote_entropy(per_learner_probs, n_classes=2):
    """modAL.disagreement.vote_entropy [external; SURVEY App. B]: entropy of vote fractions."""
    votes = vote(per_learner_probs)
    K = votes.shape[1]
    p_vote = np.zeros(shape=(votes.shape[0], n_classes))
    for c in range(n_classes):
        p_vote[:, c] = (votes == c).sum(axis=1) / K
    return _scipy_entropy(p_vote, axis=1)


def consensus_entropy(per_learner_probs):
    """modAL.disagreement.consensus_entropy [external]: entropy(transpose(predict_proba))."""
    return np.transpose(_scipy_entropy(np.transpose(committee_predict_proba(per_learner_probs))))


def kl_max_disagreement(per_learner_probs):
    """modAL.disagreement.KL_max_disagreement [external]: max_k KL(p_k || consensus)."""
    P = vote_proba(per_learner_probs)
    cons = np.mean(P, axis=1)
    K = P.shape[1]
    kl = np.zeros((P.shape[0], K))
    for k in range(K):
        kl[:, k] = _scipy_entropy(np.transpose(P[:, k, :]), qk=np.transpose(cons))
    return np.max(kl, axis=1)


MEASURES = ("uncertainty", "neg_margin", "entropy", "vote_entropy", "consensus_entropy", "kl_max")


def utility(measure, per_learner_probs):
    """Utility vector as the reference's samplers rank it (largest first)."""
    if measure == "vote_entropy":
        return vote_entropy(per_learner_probs)
    if measure == "consensus_entropy":
        return consensus_entropy(per_learner_probs)
    if measure == "kl_max":
        return kl_max_disagreement(per_learner_probs)
    # single-estimator measures act on predict_proba of the estimator handed to the sampler; for a
    # Bagging/Committee that is the mean over learners.
    if len(per_learner_probs) == 1:
        proba = widen(per_learner_probs[0])
    else:
        proba = widen(bagging_predict([widen(p) for p in per_learner_probs]))
    if measure == "uncertainty":
        return proba_uncertainty(proba)
    if measure == "neg_margin":
        return -proba_margin(proba)                         # uncertainty.py:183 ranks -margin
    if measure == "entropy":
        return proba_entropy(proba)
    raise ValueError(measure)


# --------------------------------------------------------------------------- selection
def multi_argmax(values, n_instances=1):
    """modAL.utils.selection.multi_argmax [external]: argpartition(-values, n-1)[:n] (unordered)."""
    assert n_instances <= values.shape[0]
    return np.argpartition(-values, n_instances - 1, axis=0)[:n_instances]


def topk_lowest_index(values, n_instances):
    """The build's tie rule (north_star): n largest by (value desc, index asc), in that order."""
    values = np.asarray(values)
    order = np.lexsort((np.arange(len(values)), -values.astype(np.float64)))
    return order[:n_instances]


# --------------------------------------------------------------------------- A-LINK rule
def disparity_select(p2, p1, ratio=0.25, eps=0.05, blind=False, y=None):
    """code/ALINK.py:170-201 on column-extracted scores. p2: (T, N); p1: (N,). Returns
    (sorted query indices, ACTIVE_COUNT increment). Ties inside argsort follow the build's
    lowest-index rule (numpy's default quicksort leaves them unspecified)."""
    p2 = np.asarray(p2)
    p1 = np.asarray(p1)
    T, N = p2.shape
    sets = []
    for t in range(T):
        if blind:
            idx = [j for j in range(N) if (p2[t, j] >= 0.5) != (p1[j] >= 0.5)]      # ALINK.py:177-179
        else:
            disparities = -np.absolute(p2[t] - p1)                                   # ALINK.py:181
            order = np.lexsort((np.arange(N), disparities.astype(np.float64)))
            idx = order[:int(N * ratio)]                                             # ALINK.py:183
        sets.append(set(int(j) for j in idx))
    all_noise_works = sets[0]
    for s in sets[1:]:
        all_noise_works = all_noise_works & s                                        # ALINK.py:186-188
    query, active = [], 0
    for j in sorted(all_noise_works):
        ep = p1[j]
        if ep <= 0.5 - eps or ep >= 0.5 + eps:                                       # ALINK.py:195
            active += 1                                                              # ALINK.py:198
            if y is None or ((ep >= 0.5) == (y[j] >= 0.5)):                          # ALINK.py:196-200
                query.append(j)
    return np.array(query, dtype=np.int64), active


def roc_precompute(score_matrix, masked_matrix, thresholds, roc_case):
    """utilities/ROC_precompute.py:19-63 with the Python loops replaced by their numpy meaning (float64 matrix, as
    np.loadtxt gives): returns (true_positive_rate, false_positive_rate), one value per threshold."""
    s = np.asarray(score_matrix, dtype=np.float64)
    m = np.asarray(masked_matrix)
    iu = np.triu_indices(s.shape[0], k=1, m=s.shape[1])
    sv, mv = s[iu], m[iu]
    gen_codes, imp_codes = {1: ((1,), (3,)), 2: ((2,), (4,)), 3: ((1, 2), (3, 4))}[roc_case]
    gen, imp = sv[np.isin(mv, gen_codes)], sv[np.isin(mv, imp_codes)]
    tpr = [float(np.sum(gen >= t)) / len(gen) for t in thresholds]
    fpr = [float(np.sum(imp >= t)) / len(imp) for t in thresholds]
    return np.array(tpr), np.array(fpr)


def train_step(XL, XR, y, params, acc_g, acc_d, lr=1.0, rho=0.95, eps=1e-7, act="softmax"):
    """Independent float64 restatement of one Keras batch of code/siamese.py:52-58 (finetune -> fit): forward of
    the head (:24-33), K.binary_crossentropy on the clipped activation output averaged over outputs and batch,
    binary accuracy, backprop, Adadelta (:35; Keras 2.1.2 formulas). params / acc_g / acc_d: lists of six float64
    arrays (W1,b1,W2,b2,W3,b3). Returns (loss, accuracy, new_params, new_acc_g, new_acc_d, probs)."""
    W1, b1, W2, b2, W3, b3 = [np.asarray(p, np.float64) for p in params]
    x = np.abs(np.asarray(XL, np.float64) - np.asarray(XR, np.float64))
    z1 = x @ W1 + b1
    h1 = np.maximum(z1, 0)
    z2 = h1 @ W2 + b2
    h2 = np.maximum(z2, 0)
    z3 = h2 @ W3 + b3
    y = np.asarray(y, np.float64).reshape(-1)
    nb = len(y)
    if act == "softmax":
        e = np.exp(z3 - z3.max(axis=1, keepdims=True))
        p = e / e.sum(axis=1, keepdims=True)
        t = np.stack([1 - y, y], axis=1)  # to_categorical
    else:
        p = 1 / (1 + np.exp(-z3))
        t = y[:, None]
    pc = np.clip(p, 1e-7, 1 - 1e-7)
    loss = float(np.mean(np.mean(-(t * np.log(pc) + (1 - t) * np.log(1 - pc)), axis=1)))
    acc = float(np.mean(np.mean((p > 0.5).astype(np.float64) == t, axis=1)))
    dldp = np.where((p >= 1e-7) & (p <= 1 - 1e-7), (1 - t) / (1 - pc) - t / pc, 0.0) / p.shape[1] / nb
    if act == "softmax":
        dz3 = p * (dldp - np.sum(p * dldp, axis=1, keepdims=True))
    else:
        dz3 = dldp * p * (1 - p)
    dz2 = (dz3 @ W3.T) * (h2 > 0)
    dz1 = (dz2 @ W2.T) * (h1 > 0)
    grads = [x.T @ dz1, dz1.sum(0), h1.T @ dz2, dz2.sum(0), h2.T @ dz3, dz3.sum(0)]
    new_p, new_a, new_d = [], [], []
    for P, g, a, d in zip([W1, b1, W2, b2, W3, b3], grads, acc_g, acc_d):
        a = rho * np.asarray(a, np.float64) + (1 - rho) * g * g
        u = g * np.sqrt(np.asarray(d, np.float64) + eps) / np.sqrt(a + eps)
        new_p.append(P - lr * u)
        new_a.append(a)
        new_d.append(rho * np.asarray(d, np.float64) + (1 - rho) * u * u)
    return loss, acc, new_p, new_a, new_d, p


def create_mini_batch_dfw(X_plain, X_dig):
    """code/readDFW.py:222-244, loop for loop (test infrastructure: O(N) Python)."""
    X_left, X_right, Y = [], [], []
    for i in range(len(X_plain)):
        for j in range(len(X_dig)):
            for x in X_plain[i]:
                for y in X_dig[j]:
                    X_left.append(x)
                    X_right.append(y)
                    Y.append([1] if i == j else [0])
    for i in range(len(X_dig)):
        for j in range(len(X_dig)):
            for x in X_dig[i]:
                for y in X_dig[j]:
                    X_left.append(x)
                    X_right.append(y)
                    Y.append([1] if i == j else [0])
    return [np.stack(X_left), np.stack(X_right)], np.stack(Y)


def create_mini_batch_mtp(X_plain):
    """code/readMTP.py:123-135."""
    X_left, X_right, Y = [], [], []
    for i in range(len(X_plain)):
        for j in range(len(X_plain)):
            for x in X_plain[i]:
                for y in X_plain[j]:
                    X_left.append(x)
                    X_right.append(y)
                    Y.append([1] if i == j else [0])
    return [np.stack(X_left), np.stack(X_right)], np.stack(Y)


# ---- batched differential evolution, one generation (code/differential_evolution.py:673-748) in the PARALLEL form the
# CUDA kernels use (alink_b200/csrc/de.cuh); checked against the reference solver's own output (tests/golden/ref_de.npz)
def de_mutate(pop, scale, r0, r1, cross=None):
    """best1bin trial vectors of one problem: pop (n, G) float64 -> (trial, flags of genes outside [0, 1])."""
    bprime = pop[0][None, :] + scale * (pop[r0] - pop[r1])  # :820-826
    trial = bprime if cross is None else np.where(cross.astype(bool), bprime, pop)
    return trial, (trial < 0) | (trial > 1)


def de_repair(trial, flags, rnd):
    """_ensure_constraint (:775-780): the k-th flagged gene in (member, gene) order takes rnd[k]."""
    t = trial.copy()
    t[flags] = rnd[: int(flags.sum())]  # boolean indexing enumerates row-major = (member, gene) order
    return t


def de_first(pop, e):
    """_calculate_population_energies (:660-668): the best member swaps into slot 0."""
    m = int(np.argmin(e))
    pop, e = pop.copy(), e.copy()
    e[[0, m]] = e[[m, 0]]
    pop[[0, m]] = pop[[m, 0]]
    return pop, e


def de_select(pop, e, trial, e_trial):
    """:704-714 without the sequential loop: member c takes a LOWER-energy trial; slot 0 then takes the accepted trial
    (c > 0) of lowest energy if it beats slot 0 — the first such member on ties."""
    pop, e = pop.copy(), e.copy()
    acc = e_trial < e
    pop[acc] = trial[acc]
    e[acc] = e_trial[acc]
    cand = np.nonzero(acc)[0]
    cand = cand[cand > 0]
    if len(cand):
        c = cand[np.argmin(e_trial[cand])]  # argmin: first minimum
        if e_trial[c] < e[0]:
            e[0] = e_trial[c]
            pop[0] = trial[c]
    return pop, e


# ---- A2-LINK noise-augmented pool: the counter-based stream and row arithmetic of csrc/perturb.cuh
def _u01(seed, a, b, c, d):
    """de_u01 of csrc/de.cuh on uint64 arrays (wrap-around arithmetic)."""
    with np.errstate(over="ignore"):
        u = np.uint64
        z = u(seed) ^ (np.asarray(a, u) * u(0x9E3779B97F4A7C15)) ^ (np.asarray(b, u) * u(0xBF58476D1CE4E5B9)) \
            ^ (np.asarray(c, u) * u(0x94D049BB133111EB)) ^ (np.asarray(d, u) * u(0xD6E8FEB86659FD93))
        z = z + u(0x9E3779B97F4A7C15)
        z = (z ^ (z >> u(30))) * u(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> u(27))) * u(0x94D049BB133111EB)
        z = z ^ (z >> u(31))
    return (z >> u(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def perturb_rows(table, idx, side, t, seed, pair_base=0, ncoord=40, vmax=0.25, normalize=True):
    """e' = normalize(e + delta), delta = ncoord coordinates set to U(-vmax, vmax); bit-identical to perturb_rows_kernel:
    fp32 single operations, the sum of squares in the kernel's order (lane l takes float4 chunks l, l + 32, ...; xor tree)."""
    table = np.asarray(table, np.float32)
    rows = np.arange(len(table)) if idx is None else np.asarray(idx)
    N, D = len(rows), table.shape[1]
    out = table[rows].copy()
    pair = (np.arange(N, dtype=np.uint64) + np.uint64(pair_base))
    key = np.uint64(2 * t + side)
    for j in range(ncoord):
        c = np.minimum((_u01(seed, pair, key, 2 * j, 0x5EED) * float(D)).astype(np.int64), D - 1)
        v = ((2.0 * _u01(seed, pair, key, 2 * j + 1, 0x5EED) - 1.0) * float(np.float32(vmax))).astype(np.float32)
        out[np.arange(N), c] = table[rows, c] + v          # later draws overwrite; ONE fp32 addition to the base value
    if not normalize:
        return out
    sq = out * out                                          # fp32 products
    ch = sq.reshape(N, D // 4, 4)
    nchunk = D // 4
    lanes = np.zeros((N, 32), np.float32)
    for i0 in range(0, nchunk, 32):                         # lane l: chunk i0 + l, elements x y z w in order
        blk = ch[:, i0:i0 + 32, :]
        for e in range(4):
            lanes[:, :blk.shape[1]] = lanes[:, :blk.shape[1]] + blk[:, :, e]
    for o in (16, 8, 4, 2, 1):                              # ss += shfl_xor(ss, o)
        lanes = lanes + lanes[:, np.arange(32) ^ o]
    ss = lanes[:, 0]
    inv = (np.float32(1.0) / np.sqrt(ss)).astype(np.float32)
    return (out * inv[:, None]).astype(np.float32)
