"""ctypes binding of oracle/libalink_oracle.so. TEST INFRASTRUCTURE ONLY — see alink_oracle.c header.

Imported by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs; never by alink_b200/.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libalink_oracle.so")

MEASURE_ID = {"uncertainty": 0, "neg_margin": 1, "entropy": 2, "vote_entropy": 3,
              "consensus_entropy": 4, "kl_max": 5}
ACT_ID = {"softmax": 0, "sigmoid": 1}


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "alink_oracle.c")
    hdr = os.path.join(_HERE, "..", "include", "alink_math.h")
    stale = (not os.path.exists(_SO)) or any(
        os.path.exists(f) and os.path.getmtime(f) > os.path.getmtime(_SO) for f in (src, hdr))
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "-B", "libalink_oracle.so"],
                              stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        fpp = C.POINTER(C.c_void_p)
        L.alink_oracle_score.restype = C.c_int
        L.alink_oracle_score.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                         C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                         fpp, fpp, fpp, fpp, fpp, fpp,
                                         C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.alink_oracle_utility.restype = C.c_int
        L.alink_oracle_utility.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.alink_oracle_topk.restype = C.c_int
        L.alink_oracle_topk.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
        L.alink_oracle_topk_keys.restype = C.c_int
        L.alink_oracle_topk_keys.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p]
        L.alink_oracle_disparity_select.restype = C.c_int
        L.alink_oracle_disparity_select.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_double,
                                                    C.c_double, C.c_int, C.c_void_p, C.c_void_p,
                                                    C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.alink_oracle_expf.restype = C.c_float
        L.alink_oracle_expf.argtypes = [C.c_float]
        L.alink_oracle_logf.restype = C.c_float
        L.alink_oracle_logf.argtypes = [C.c_float]
        L.alink_oracle_pack_key.restype = C.c_uint64
        L.alink_oracle_pack_key.argtypes = [C.c_float, C.c_uint64]
        L.alink_oracle_set_threads.argtypes = [C.c_int]
        L.alink_oracle_get_threads.restype = C.c_int
        _lib = L
    return _lib


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _pp(arrs):
    arr = (C.c_void_p * len(arrs))(*[a.ctypes.data for a in arrs])
    return C.cast(arr, C.POINTER(C.c_void_p)), arr


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def score(table_a, table_b, ia, ib, heads, feat_mode=0, measure="uncertainty",
          want_per_learner=True, want_consensus=True, want_util=True, threads=0):
    """Returns dict(per_learner (N,K,C), consensus (N,C), util (N,)) computed by the C oracle."""
    L = lib()
    L.alink_oracle_set_threads(int(threads))
    ta, tb = _f32(table_a), _f32(table_b)
    D = ta.shape[1]
    if ia is None:
        N = ta.shape[0]
        ia_ = ib_ = None
    else:
        ia_ = np.ascontiguousarray(ia, dtype=np.int32)
        ib_ = np.ascontiguousarray(ib, dtype=np.int32)
        N = len(ia_)
    K, Cc = heads.K, heads.C
    keep = []
    ptrs = []
    for lst in (heads.W1, heads.b1, heads.W2, heads.b2, heads.W3, heads.b3):
        arrs = [_f32(a) for a in lst]
        p, holder = _pp(arrs)
        keep.append((arrs, holder))
        ptrs.append(p)
    per = np.empty((N, K, Cc), np.float32) if want_per_learner else None
    cons = np.empty((N, Cc), np.float32) if want_consensus else None
    util = np.empty((N,), np.float32) if want_util else None
    rc = L.alink_oracle_score(_ptr(ta), _ptr(tb), _ptr(ia_), _ptr(ib_), N, D, int(feat_mode), K,
                              heads.H1, heads.H2, Cc, ACT_ID[heads.act], *ptrs,
                              _ptr(per), _ptr(cons), MEASURE_ID[measure], _ptr(util))
    if rc != 0:
        raise ValueError("alink_oracle_score rc=%d" % rc)
    return dict(per_learner=per, consensus=cons, util=util)


def utility(per_learner, measure):
    L = lib()
    per = _f32(per_learner)
    N, K, Cc = per.shape
    util = np.empty((N,), np.float32)
    cons = np.empty((N, Cc), np.float32)
    rc = L.alink_oracle_utility(_ptr(per), N, K, Cc, MEASURE_ID[measure], _ptr(util), _ptr(cons))
    if rc != 0:
        raise ValueError("alink_oracle_utility rc=%d" % rc)
    return util, cons


def topk(util, B, base=0):
    L = lib()
    u = _f32(util)
    idx = np.empty((B,), np.int64)
    uo = np.empty((B,), np.float32)
    keys = np.empty((B,), np.uint64)
    rc = L.alink_oracle_topk(_ptr(u), len(u), int(base), int(B), _ptr(idx), _ptr(uo), _ptr(keys))
    if rc != 0:
        raise ValueError("alink_oracle_topk rc=%d" % rc)
    return idx, uo, keys


def topk_keys(keys, B):
    L = lib()
    k = np.ascontiguousarray(keys, dtype=np.uint64)
    out = np.empty((B,), np.uint64)
    rc = L.alink_oracle_topk_keys(_ptr(k), len(k), int(B), _ptr(out))
    if rc != 0:
        raise ValueError("alink_oracle_topk_keys rc=%d" % rc)
    return out


def disparity_select(p2, p1, ratio=0.25, eps=0.05, blind=False, y=None):
    L = lib()
    p2 = _f32(p2)
    p1 = _f32(p1)
    T, N = p2.shape
    yy = None if y is None else _f32(y).reshape(-1)
    idx = np.empty((max(N, 1),), np.int64)
    n_out = C.c_int64(0)
    active = C.c_int64(0)
    rc = L.alink_oracle_disparity_select(_ptr(p2), _ptr(p1), N, T, float(ratio), float(eps), int(bool(blind)),
                                         _ptr(yy), _ptr(idx), C.byref(n_out), C.byref(active))
    if rc != 0:
        raise ValueError("alink_oracle_disparity_select rc=%d" % rc)
    return idx[:n_out.value].copy(), int(active.value)


def expf(x):
    return float(lib().alink_oracle_expf(float(x)))


def logf(x):
    return float(lib().alink_oracle_logf(float(x)))


def pack_key(u, idx):
    return int(lib().alink_oracle_pack_key(float(u), int(idx)))


def roc_counts(scores, mask, thresholds, roc_case, row0=0):
    """(tp, fp, n_genuine, n_impostor) — utilities/ROC_precompute.py:19-63 on a block of the score matrix."""
    L = lib()
    L.alink_oracle_roc_counts.restype = C.c_int
    L.alink_oracle_roc_counts.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int, C.c_void_p,
                                          C.c_int, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    s = _f32(scores)
    m = np.ascontiguousarray(mask, dtype=np.int32)
    thr = np.ascontiguousarray(thresholds, dtype=np.float64).reshape(-1)
    tp, fp = np.zeros(len(thr), np.uint64), np.zeros(len(thr), np.uint64)
    ng, ni = C.c_uint64(0), C.c_uint64(0)
    rc = L.alink_oracle_roc_counts(_ptr(s), _ptr(m), s.shape[0], s.shape[1], int(row0), int(roc_case), _ptr(thr),
                                   len(thr), _ptr(tp), _ptr(fp), C.byref(ng), C.byref(ni))
    if rc != 0:
        raise ValueError("alink_oracle_roc_counts rc=%d" % rc)
    return tp, fp, int(ng.value), int(ni.value)


def train_step(table_a, table_b, ia, ib, y, params, acc_g, acc_d, D, H1, H2, Cc, feat_mode=0, lr=1.0, rho=0.95,
               eps=1e-7, apply=True, want_probs=False):
    """One finetune minibatch step of ONE learner (code/siamese.py:52-58 per batch). params / acc_g / acc_d: lists of
    six float32 arrays (W1,b1,W2,b2,W3,b3), UPDATED IN PLACE when apply. Returns (loss, accuracy[, probs])."""
    L = lib()
    L.alink_oracle_train_step.restype = C.c_int
    L.alink_oracle_train_step.argtypes = [C.c_void_p] * 5 + [C.c_int] * 6 + [C.POINTER(C.c_void_p)] * 3 + \
        [C.c_float] * 3 + [C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float), C.c_void_p]
    ta, tb = _f32(table_a), _f32(table_b)
    yy = _f32(y).reshape(-1)
    nb = len(yy)
    ia_ = None if ia is None else np.ascontiguousarray(ia, dtype=np.int32)
    ib_ = None if ib is None else np.ascontiguousarray(ib, dtype=np.int32)
    for lst in (params, acc_g, acc_d):
        for a in lst:
            assert a.dtype == np.float32 and a.flags["C_CONTIGUOUS"]

    def pp(lst):
        return (C.c_void_p * 6)(*[a.ctypes.data for a in lst])
    loss, acc = C.c_float(0), C.c_float(0)
    probs = np.empty((nb, Cc), np.float32) if want_probs else None
    rc = L.alink_oracle_train_step(_ptr(ta), _ptr(tb), _ptr(ia_), _ptr(ib_), _ptr(yy), nb, D, feat_mode, H1, H2, Cc,
                                   pp(params), pp(acc_g), pp(acc_d), lr, rho, eps, int(bool(apply)),
                                   C.byref(loss), C.byref(acc), _ptr(probs))
    if rc != 0:
        raise ValueError("alink_oracle_train_step rc=%d" % rc)
    return (loss.value, acc.value, probs) if want_probs else (loss.value, acc.value)
