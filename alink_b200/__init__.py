"""alink_b200 — B200-native drop-in for A-LINK's active-learning query-selection step.

Module names mirror /root/reference/code: siamese, siamese3, committee, keras_wrapper, uncertainty,
base, learners (+ disagreement = modAL.disagreement, alink = the paper's disparity rule). The compute
lives in csrc/ (CUDA, sm_100a) behind the C ABI of include/alink_b200.h; `_ffi` binds it with ctypes.
"""
from . import _ffi, engine  # noqa: F401
from .engine import PairPool, set_compute_mode  # noqa: F401

__all__ = ["PairPool", "set_compute_mode"]
