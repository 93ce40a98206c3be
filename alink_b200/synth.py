"""Synthetic pools for the query-selection path.

The reference ships no data, weights or fixtures (datasets are DFW / Multi-PIE downloads,
/root/reference/code/arcface_prepreq.sh:7-20), so every test and benchmark input is generated
here, deterministically, with numpy `default_rng(PCG64)`; base seed 42 is the reference's own
(`tf.set_random_seed(42)`, /root/reference/code/ALINK.py:19).

Shapes follow the reference's feature extractors:
  * ArcFace r100, 512-d, L2-normalised ........ /root/reference/code/face_model.py:86-93
  * VGGFace-ResNet50 avg_pool, 2048-d, >= 0 .... /root/reference/code/siamese.py:203-216
Heads follow Keras-2.1.2 Dense defaults (glorot_uniform kernel) for
/root/reference/code/siamese.py:24-33 and code/siamese3.py:18-26, with the last layer's gain
calibrated so that the logit margin has std 2.0 (otherwise every pair ties at p ~ 0.5).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List

import numpy as np

BASE_SEED = 42


def make_table(M: int, D: int, n_ident: int, kind: str = "arcface", seed: int = BASE_SEED) -> np.ndarray:
    """(M, D) fp32 embedding table; row m belongs to identity m // (M // n_ident)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    per = max(M // n_ident, 1)
    ident = np.minimum(np.arange(M) // per, n_ident - 1)
    if kind == "arcface":
        c = rng.standard_normal((n_ident, D), dtype=np.float32)
        c /= np.linalg.norm(c, axis=1, keepdims=True)
        n = rng.standard_normal((M, D), dtype=np.float32) * np.float32(1.0 / np.sqrt(D))
        e = c[ident] + np.float32(0.6) * n
        e /= np.linalg.norm(e, axis=1, keepdims=True)
    elif kind == "resnet50":
        c = rng.standard_normal((n_ident, D), dtype=np.float32)
        n = rng.standard_normal((M, D), dtype=np.float32)
        e = np.maximum(c[ident] + np.float32(0.5) * n, 0)
    else:
        raise ValueError(kind)
    return np.ascontiguousarray(e, dtype=np.float32)


def make_pairs(N: int, M: int, n_ident: int, seed: int = BASE_SEED + 1):
    """int32 (ia, ib): ia uniform; ib genuine (same identity) w.p. 0.5 else uniform. Pair index = position."""
    rng = np.random.Generator(np.random.PCG64(seed))
    per = max(M // n_ident, 1)
    ia = rng.integers(0, M, size=N, dtype=np.int64)
    genuine = rng.random(N) < 0.5
    ident = np.minimum(ia // per, n_ident - 1)
    same = np.minimum(ident * per + rng.integers(0, per, size=N, dtype=np.int64), M - 1)
    other = rng.integers(0, M, size=N, dtype=np.int64)
    ib = np.where(genuine, same, other)
    return ia.astype(np.int32), ib.astype(np.int32), genuine


@dataclass
class Heads:
    """K siamese heads in Keras layout: W (in, out) row-major fp32, b (out,)."""
    K: int
    D_in: int
    H1: int
    H2: int
    C: int
    act: str  # "softmax" (C=2, siamese.py) | "sigmoid" (C=1, siamese3.py)
    W1: List[np.ndarray] = field(default_factory=list)
    b1: List[np.ndarray] = field(default_factory=list)
    W2: List[np.ndarray] = field(default_factory=list)
    b2: List[np.ndarray] = field(default_factory=list)
    W3: List[np.ndarray] = field(default_factory=list)
    b3: List[np.ndarray] = field(default_factory=list)

    def learner(self, k: int):
        return self.W1[k], self.b1[k], self.W2[k], self.b2[k], self.W3[k], self.b3[k]


def _glorot(rng, fan_in, fan_out):
    lim = np.sqrt(6.0 / (fan_in + fan_out))
    return rng.uniform(-lim, lim, size=(fan_in, fan_out)).astype(np.float32)


def make_heads(K: int, D_in: int, calib_x: np.ndarray, H1: int = 512, H2: int = 64, act: str = "softmax",
               seed0: int = 1000, margin_std: float = 2.0) -> Heads:
    """Glorot heads, seed 1000+k per learner; W3 scaled so std(z1-z0) (or std(z) for sigmoid) == margin_std
    on the calibration features `calib_x` ((n, D_in) fp32, the first 4096 pairs' features)."""
    C = 2 if act == "softmax" else 1
    hs = Heads(K, D_in, H1, H2, C, act)
    calib_x = np.asarray(calib_x, dtype=np.float32)
    for k in range(K):
        rng = np.random.Generator(np.random.PCG64(seed0 + k))
        W1 = _glorot(rng, D_in, H1)
        b1 = rng.uniform(-0.05, 0.05, size=H1).astype(np.float32)
        W2 = _glorot(rng, H1, H2)
        b2 = rng.uniform(-0.05, 0.05, size=H2).astype(np.float32)
        W3 = _glorot(rng, H2, C)
        b3 = rng.uniform(-0.05, 0.05, size=C).astype(np.float32)
        h = np.maximum(calib_x @ W1 + b1, 0)
        h = np.maximum(h @ W2 + b2, 0)
        z = h @ W3
        m = (z[:, 1] - z[:, 0]) if C == 2 else z[:, 0]
        sd = float(np.std(m - m.mean()))
        if sd > 0:
            W3 = (W3 * np.float32(margin_std / sd)).astype(np.float32)
            # centre the margin so both classes occur
            z = h @ W3
            m = (z[:, 1] - z[:, 0]) if C == 2 else z[:, 0]
            if C == 2:
                b3 = b3 + np.array([0.5, -0.5], dtype=np.float32) * np.float32(m.mean())
            else:
                b3 = b3 - np.float32(m.mean())
        for lst, a in ((hs.W1, W1), (hs.b1, b1), (hs.W2, W2), (hs.b2, b2), (hs.W3, W3), (hs.b3, b3)):
            lst.append(np.ascontiguousarray(a, dtype=np.float32))
    return hs


def pair_features(table_a, table_b, ia, ib, feat_mode: int = 0) -> np.ndarray:
    a = table_a[ia]
    b = table_b[ib]
    x = np.abs(a - b)
    if feat_mode:
        x = np.concatenate([x, a * b], axis=1)
    return x.astype(np.float32)


@dataclass
class Pool:
    name: str
    table: np.ndarray
    ia: np.ndarray
    ib: np.ndarray
    heads: Heads
    B: int
    measure: str
    genuine: np.ndarray

    @property
    def N(self):
        return len(self.ia)

    @property
    def D(self):
        return self.table.shape[1]

    @property
    def M(self):
        return self.table.shape[0]


# BASELINE.json configs, 1-based ordinal as in SURVEY.md §8d
CONFIGS = {
    1: dict(N=2_000, D=512, M=1_000, I=250, K=3, B=100, kind="arcface", act="softmax"),
    2: dict(N=1_000_000, D=512, M=200_000, I=50_000, K=5, B=1_000, kind="arcface", act="softmax"),
    3: dict(N=10_000_000, D=2048, M=500_000, I=125_000, K=5, B=1_000, kind="resnet50", act="softmax"),
}


def make_pool(cfg: int = 1, N: int | None = None, M: int | None = None, K: int | None = None,
              act: str | None = None, measure: str = "uncertainty", feat_mode: int = 0,
              H1: int = 512, H2: int = 64, D: int | None = None, kind: str | None = None,
              B: int | None = None) -> Pool:
    c = dict(CONFIGS[cfg])
    if M is not None:
        c["I"] = max(int(c["I"] * M / c["M"]), 1)
        c["M"] = M
    for k_, v in (("N", N), ("K", K), ("act", act), ("D", D), ("kind", kind), ("B", B)):
        if v is not None:
            c[k_] = v
    seed = BASE_SEED + cfg
    table = make_table(c["M"], c["D"], c["I"], c["kind"], seed)
    ia, ib, genuine = make_pairs(c["N"], c["M"], c["I"], seed + 100)
    n_cal = min(4096, c["N"])
    calib = pair_features(table, table, ia[:n_cal], ib[:n_cal], feat_mode)
    D_in = c["D"] * (2 if feat_mode else 1)
    heads = make_heads(c["K"], D_in, calib, H1=H1, H2=H2, act=c["act"])
    return Pool(f"cfg{cfg}", table, ia, ib, heads, min(c["B"], c["N"]), measure, genuine)
