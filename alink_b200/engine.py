"""engine.py — host-side glue between the drop-in classes and the C ABI.

Responsibilities (all thin; the arithmetic lives in csrc/):
  * map a list of SiameseNetwork objects (a Bagging / Committee / single estimator) onto one of the
    context's head groups, re-uploading weights only when they changed;
  * turn the reference's `[X_left, X_right]` calling convention (code/ALINK.py:160, code/existing_al.py:108)
    or an index-driven pool (`PairPool`) into `alink_score_*` / `alink_query` calls.
"""
from __future__ import annotations

import itertools
import weakref
from dataclasses import dataclass
from typing import List, Optional, Sequence

import numpy as np

from . import _ffi

_COMPUTE = {"mode": _ffi.FP32_EXACT}


def set_compute_mode(mode: str) -> None:
    """'fp32' (bit-exact with the oracle, CUDA cores), 'bf16' (tcgen05 tensor cores) or 'refine' (queries: tensor-core
    filter + bit-exact re-score, selection identical to 'fp32'; score calls stay exact)."""
    _COMPUTE["mode"] = _ffi.COMPUTE_ID[mode]
    _GroupCache.invalidate_all()


def get_compute_mode() -> int:
    return _COMPUTE["mode"]


_UIDS = itertools.count(1)  # process-unique tokens: id() of a freed object is reused by CPython


class HeadSpec:
    """Weights of one learner in Keras layout (what SiameseNetwork owns)."""

    def __init__(self, D_in, H1, H2, C, act, feat_mode=0):
        self.uid = next(_UIDS)
        self.D_in, self.H1, self.H2, self.C, self.act, self.feat_mode = D_in, H1, H2, C, act, feat_mode
        self.W1 = self.b1 = self.W2 = self.b2 = self.W3 = self.b3 = None
        self.version = 0

    def set(self, W1, b1, W2, b2, W3, b3):
        f = lambda a, shp: np.ascontiguousarray(np.asarray(a, dtype=np.float32).reshape(shp))
        self.W1, self.b1 = f(W1, (self.D_in, self.H1)), f(b1, (self.H1,))
        self.W2, self.b2 = f(W2, (self.H1, self.H2)), f(b2, (self.H2,))
        self.W3, self.b3 = f(W3, (self.H2, self.C)), f(b3, (self.C,))
        self.version += 1

    @property
    def ready(self):
        return self.W1 is not None


@dataclass
class _Stack:
    K: int
    D_in: int
    H1: int
    H2: int
    C: int
    act: str
    W1: list
    b1: list
    W2: list
    b2: list
    W3: list
    b3: list


class _GroupCache:
    """LRU of head groups per context: key = identities + versions of the learners' weights."""
    _by_ctx = {}

    @classmethod
    def invalidate_all(cls):
        cls._by_ctx.clear()

    @staticmethod
    def key_of(specs: Sequence[HeadSpec]):
        """(uid, version) per learner + the compute mode: uids are never reused, unlike id()."""
        return tuple((s.uid, s.version) for s in specs) + (_COMPUTE["mode"],)

    @classmethod
    def group_for(cls, ctx: _ffi.Context, specs: Sequence[HeadSpec]) -> int:
        ngroups = 4
        st = cls._by_ctx.setdefault(_ctx_uid(ctx), {"keys": [None] * ngroups, "clock": 0, "used": [0] * ngroups})
        key = cls.key_of(specs)
        st["clock"] += 1
        if key in st["keys"]:
            g = st["keys"].index(key)
            st["used"][g] = st["clock"]
            return g
        g = int(np.argmin(st["used"]))
        s0 = specs[0]
        for s in specs:
            if not s.ready:
                raise _ffi.AlinkError("SiameseNetwork has no weights: call set_weights()/maybeLoadFromMemory() first")
            if (s.D_in, s.H1, s.H2, s.C, s.act, s.feat_mode) != (s0.D_in, s0.H1, s0.H2, s0.C, s0.act, s0.feat_mode):
                raise _ffi.AlinkError("all learners of a committee must share one head architecture")
        stack = _Stack(len(specs), s0.D_in, s0.H1, s0.H2, s0.C, s0.act,
                       [s.W1 for s in specs], [s.b1 for s in specs], [s.W2 for s in specs],
                       [s.b2 for s in specs], [s.W3 for s in specs], [s.b3 for s in specs])
        ctx.set_heads(g, stack, feat_mode=s0.feat_mode, compute=_COMPUTE["mode"])
        st["keys"][g] = key
        st["used"][g] = st["clock"]
        return g


def _ctx_uid(ctx) -> int:
    """A process-unique token per context object (dropped from the cache when the context is collected)."""
    uid = getattr(ctx, "_alink_uid", None)
    if uid is None:
        uid = next(_UIDS)
        ctx._alink_uid = uid
        weakref.finalize(ctx, _GroupCache._by_ctx.pop, uid, None)
    return uid


def _as_pair(X):
    """The reference passes X = [X_left, X_right] (list of two (N,D) arrays)."""
    if isinstance(X, PairPool):
        return X
    if not (isinstance(X, (list, tuple)) and len(X) == 2):
        raise ValueError("expected [X_left, X_right] or a PairPool")
    XL = np.ascontiguousarray(X[0], dtype=np.float32)
    XR = np.ascontiguousarray(X[1], dtype=np.float32)
    if XL.ndim != 2 or XL.shape != XR.shape:
        raise ValueError("X_left and X_right must both be (N, D)")
    return XL, XR


class PairPool:
    """Index-driven pool: pair n = (table_a[ia[n]], table_b[ib[n]]).

    The B200-native representation of the reference's materialising loops
    (code/readDFW.py:222-244, utilities/generateMatrixDFW.py:28-33). Tables / indices may be numpy
    arrays (copied to the GPU once, here) or torch CUDA tensors (borrowed)."""

    # Table slots 0 .. POOL_SLOTS-1 belong to pools; MAX_SLOTS-3 is reserved for testAccuracy / score_matrix and
    # MAX_SLOTS-2 / -1 for the dense [X_left, X_right] staging (alink_score_dense). A slot is owned by ONE live pool:
    # asking for more pools than there are slots raises instead of silently overwriting a live pool's table.
    POOL_SLOTS = _ffi.MAX_SLOTS - 3
    _owners = {}  # context uid -> [weakref to the owning PairPool or None] * POOL_SLOTS

    @classmethod
    def _claim(cls, ctx, pool, n):
        own = cls._owners.setdefault(_ctx_uid(ctx), [None] * cls.POOL_SLOTS)
        free = [i for i, w in enumerate(own) if w is None or w() is None]
        if len(free) < n:
            raise _ffi.AlinkError(f"no free embedding-table slot: {cls.POOL_SLOTS} pool slots per context are all owned "
                                  "by live PairPools (close() or drop one first)")
        for i in free[:n]:
            own[i] = weakref.ref(pool)
        return free[:n]

    def close(self):
        """Give the pool's table slots back (also happens when the pool is garbage-collected)."""
        own = PairPool._owners.get(getattr(self.ctx, "_alink_uid", None))
        if own is not None:
            for i, w in enumerate(own):
                if w is not None and w() is self:
                    own[i] = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __init__(self, table_a, ia, ib, table_b=None, ctx: Optional[_ffi.Context] = None, slots=None):
        self.ctx = ctx or _ffi.default_context()
        if slots is None:
            got = PairPool._claim(self.ctx, self, 1 if table_b is None else 2)
            slots = (got[0], got[-1])
        self.slot_a, self.slot_b = slots
        self.ctx.set_table(self.slot_a, table_a)
        if table_b is None:
            self.slot_b = self.slot_a
        else:
            self.ctx.set_table(self.slot_b, table_b)
        self.table_a, self.table_b = table_a, (table_a if table_b is None else table_b)
        self.ia, self.ib = ia, ib
        for name, idx, t in (("ia", ia, self.table_a), ("ib", ib, self.table_b)):
            if isinstance(idx, np.ndarray) and idx.size:  # once per pool; device tensors: Context(check_indices=True)
                lo, hi = int(idx.min()), int(idx.max())
                if lo < 0 or hi >= int(t.shape[0]):
                    raise IndexError(f"{name} holds index {lo if lo < 0 else hi}, the table has {int(t.shape[0])} rows")
        self.N = int(len(ia)) if ia is not None else int(table_a.shape[0])
        self.D = int(table_a.shape[1])

    def __len__(self):
        return self.N

    def left_rows(self, idx):
        """Materialise X_left rows for the selected pairs (what the reference's samplers return)."""
        rows = np.asarray(self.ia)[idx] if self.ia is not None and not _ffi._is_torch(self.ia) else idx
        if _ffi._is_torch(self.ia):
            rows = self.ia[idx].cpu().numpy()
        t = self.table_a
        return t[rows].cpu().numpy() if _ffi._is_torch(t) else t[rows]


def score(specs: Sequence[HeadSpec], X, want_per=True, want_cons=True, ctx=None):
    """per_learner (N,K,C), consensus (N,C) fp32 for the learners in `specs`."""
    ctx = ctx or (X.ctx if isinstance(X, PairPool) else _ffi.default_context())
    g = _GroupCache.group_for(ctx, specs)
    K, Cc = len(specs), specs[0].C
    if isinstance(X, PairPool):
        return ctx.score_indexed(g, X.slot_a, X.slot_b, X.ia, X.ib, X.N, K, Cc, want_per, want_cons)
    XL, XR = _as_pair(X)
    if len(XL) == 0:
        return (np.zeros((0, K, Cc), np.float32) if want_per else None,
                np.zeros((0, Cc), np.float32) if want_cons else None)
    return ctx.score_dense(g, XL, XR, K, Cc, want_per, want_cons)


def fused_query(specs: Sequence[HeadSpec], X, measure: str, n_instances: int, base: int = 0, ctx=None,
                timing: bool = False):
    """score -> K-learner reduce -> utility -> top-B in one C-ABI call. Returns (idx, util, keys[, timing])."""
    ctx = ctx or (X.ctx if isinstance(X, PairPool) else _ffi.default_context())
    g = _GroupCache.group_for(ctx, specs)
    if isinstance(X, PairPool):
        return ctx.query(g, X.slot_a, X.slot_b, X.ia, X.ib, X.N, measure, n_instances, base, timing)
    XL, XR = _as_pair(X)
    ctx.set_table(_ffi.MAX_SLOTS - 2, XL)
    ctx.set_table(_ffi.MAX_SLOTS - 1, XR)
    return ctx.query(g, _ffi.MAX_SLOTS - 2, _ffi.MAX_SLOTS - 1, None, None, len(XL), measure, n_instances, base,
                     timing)


def specs_of(estimator) -> Optional[List[HeadSpec]]:
    """HeadSpecs behind an estimator-like object, or None if it is not one of ours."""
    get = getattr(estimator, "_alink_specs", None)
    return get() if callable(get) else None
