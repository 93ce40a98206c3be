"""ctypes binding of libalink_b200.so (C ABI: include/alink_b200.h).

This is the only bridge between the Python host code and the CUDA kernels. There is no CPU
fallback: if the shared library is missing or no CUDA device is visible, every entry point
raises — loudly — instead of computing on the host.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# ALINK_LIB_DEBUG=1 loads the debug build (same ABI; role-delay injection and the multi-record watchdog of the
# tensor-core head: tools/stress_roles.py, tests/test_gpu_stress.py)
SO_PATH = os.path.join(_HERE, "csrc", "libalink_b200_dbg.so" if os.environ.get("ALINK_LIB_DEBUG") == "1"
                       else "libalink_b200.so")
if os.environ.get("ALINK_LIB_PATH"):  # A/B runs of two builds inside one GPU session (tools/)
    SO_PATH = os.environ["ALINK_LIB_PATH"]

HOST, DEVICE = 0, 1
F32, BF16 = 0, 1
FEAT_ABSDIFF, FEAT_ABSDIFF_HADAMARD = 0, 1
FP32_EXACT, BF16_TC, FP32_REFINE = 0, 1, 2
E_CUDA = -2
ACT_ID = {"softmax": 0, "sigmoid": 1}
MEASURE_ID = {"uncertainty": 0, "neg_margin": 1, "entropy": 2, "vote_entropy": 3,
              "consensus_entropy": 4, "kl_max": 5}
COMPUTE_ID = {"fp32": FP32_EXACT, "fp32_exact": FP32_EXACT, "bf16": BF16_TC, "bf16_tc": BF16_TC,
              "refine": FP32_REFINE, "fp32_refine": FP32_REFINE}
MAX_SLOTS = 8

# every symbol include/alink_b200.h declares (tests/test_abi.py checks the .so exports them all)
SYMBOLS = (
    "alink_abi_version", "alink_create", "alink_destroy", "alink_last_error", "alink_sync",
    "alink_set_stream", "alink_set_reserved_sms", "alink_launch_count", "alink_watchdog_record", "alink_host_alloc", "alink_host_free", "alink_set_heads", "alink_set_table", "alink_set_table_async",
    "alink_score_indexed", "alink_score_dense", "alink_utility", "alink_utility_from_proba",
    "alink_select_topk", "alink_topk_values", "alink_topk_values_noalloc", "alink_topk_keys", "alink_query", "alink_disparity_select",
    "alink_set_refine_eps", "alink_refine_stats", "alink_refine_detail", "alink_score_allpairs", "alink_query_allpairs",
    "alink_disparity_hist", "alink_disparity_select_sharded", "alink_roc_counts",
    "alink_train_step", "alink_train_epoch", "alink_get_train_state", "alink_set_train_state",
    "alink_set_refine_eps2", "alink_refine_report_get",
    "alink_perturb_rows", "alink_get_table_rows", "alink_score_column", "alink_min_disparity",
    "alink_comm_unique_id", "alink_comm_init", "alink_comm_destroy", "alink_query_sharded", "alink_comm_push_state", "alink_disparity_select_comm",
    "alink_de_init", "alink_de_mutate", "alink_de_repair", "alink_de_evaluate", "alink_de_select", "alink_de_get",
)


class AlinkError(RuntimeError):
    pass


class Timing(C.Structure):
    _fields_ = [("ms_h2d", C.c_float), ("ms_score", C.c_float), ("ms_utility", C.c_float),
                ("ms_select", C.c_float), ("ms_d2h", C.c_float), ("ms_total", C.c_float),
                ("launches", C.c_int64)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class RefineReport(C.Structure):
    _fields_ = [("version", C.c_int32), ("filter_f16", C.c_int32), ("rounds", C.c_int32), ("fallback", C.c_int32),
                ("n_stage2", C.c_int64), ("n_exact", C.c_int64), ("n_sample", C.c_int64), ("n_sample_exact", C.c_int64),
                ("eps1_used", C.c_double), ("eps1_apriori", C.c_double), ("eps2_used", C.c_double), ("eps1_worst", C.c_double),
                ("err1_max", C.c_float), ("err1_rms", C.c_float), ("err2_max", C.c_float), ("err2_rms", C.c_float)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


_lib = None


def load():
    """Load the product library; raise if it has not been built (see __graft_entry__.build())."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise AlinkError(
            f"{SO_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C alink_b200/csrc`). There is no CPU fallback.")
    L = C.CDLL(SO_PATH)
    vp, i32, i64, u64, dbl = C.c_void_p, C.c_int, C.c_int64, C.c_uint64, C.c_double
    pp = C.POINTER(C.c_void_p)
    L.alink_abi_version.restype = i32
    L.alink_create.restype = i32
    L.alink_create.argtypes = [C.POINTER(vp), i32, C.c_uint]
    L.alink_destroy.restype = i32
    L.alink_destroy.argtypes = [vp]
    L.alink_last_error.restype = C.c_char_p
    L.alink_last_error.argtypes = [vp]
    L.alink_sync.restype = i32
    L.alink_sync.argtypes = [vp]
    L.alink_set_stream.restype = i32
    L.alink_set_stream.argtypes = [vp, vp]
    L.alink_launch_count.restype = i64
    L.alink_launch_count.argtypes = [vp]
    L.alink_host_alloc.restype = i32
    L.alink_host_alloc.argtypes = [C.POINTER(vp), u64]
    L.alink_host_free.restype = i32
    L.alink_host_free.argtypes = [vp]
    L.alink_set_heads.restype = i32
    L.alink_set_heads.argtypes = [vp, i32, i32, i32, i32, i32, i32, i32, i32, pp, pp, pp, pp, pp, pp, i32]
    L.alink_set_table.restype = i32
    L.alink_set_table.argtypes = [vp, i32, vp, i64, i32, i32, i32]
    L.alink_set_table_async.restype = i32
    L.alink_set_table_async.argtypes = [vp, i32, vp, i64, i32, i32]
    L.alink_score_indexed.restype = i32
    L.alink_score_indexed.argtypes = [vp, i32, i32, i32, vp, vp, i64, i32, vp, vp, i32]
    L.alink_score_dense.restype = i32
    L.alink_score_dense.argtypes = [vp, i32, vp, vp, i64, i32, i32, vp, vp, i32]
    L.alink_utility.restype = i32
    L.alink_utility.argtypes = [vp, i32, vp, i32]
    L.alink_utility_from_proba.restype = i32
    L.alink_utility_from_proba.argtypes = [vp, vp, i64, i32, i32, i32, i32, vp, vp, i32]
    L.alink_select_topk.restype = i32
    L.alink_select_topk.argtypes = [vp, i64, i64, vp, vp, vp, i32]
    L.alink_topk_values.restype = i32
    L.alink_topk_values.argtypes = [vp, vp, i64, i64, i64, i32, vp, vp, vp, i32]
    L.alink_topk_values_noalloc.restype = i32
    L.alink_topk_values_noalloc.argtypes = [vp, vp, i64, i64, i64, vp, vp, vp]
    L.alink_topk_keys.restype = i32
    L.alink_topk_keys.argtypes = [vp, vp, i64, i64, i32, vp, i32]
    L.alink_query.restype = i32
    L.alink_query.argtypes = [vp, i32, i32, i32, vp, vp, i64, i32, i32, i64, i64, vp, vp, vp, i32,
                              C.POINTER(Timing)]
    L.alink_disparity_select.restype = i32
    L.alink_disparity_select.argtypes = [vp, vp, vp, i64, i32, dbl, dbl, i32, vp, i32, vp, i32,
                                         C.POINTER(i64), C.POINTER(i64)]
    L.alink_score_allpairs.restype = i32
    L.alink_score_allpairs.argtypes = [vp, i32, i32, i32, i64, i64, vp, vp, i32]
    L.alink_query_allpairs.restype = i32
    L.alink_query_allpairs.argtypes = [vp, i32, i32, i32, i64, i64, i32, i64, vp, vp, vp, i32, C.POINTER(Timing)]
    L.alink_set_refine_eps.restype = i32
    L.alink_set_refine_eps.argtypes = [vp, dbl]
    L.alink_refine_stats.restype = i32
    L.alink_refine_stats.argtypes = [vp, C.POINTER(i64), C.POINTER(C.c_float)]
    L.alink_train_epoch.restype = i32
    L.alink_train_epoch.argtypes = [vp, i32, i32, i32, i32, vp, vp, vp, i64, i32, i32, dbl, dbl, dbl, i32, vp]
    L.alink_train_step.restype = i32
    L.alink_train_step.argtypes = [vp, i32, i32, i32, i32, vp, vp, vp, i32, i32, dbl, dbl, dbl, i32,
                                   C.POINTER(C.c_float), C.POINTER(C.c_float), vp]
    L.alink_get_train_state.restype = i32
    L.alink_get_train_state.argtypes = [vp, i32, i32, i32, C.POINTER(vp)]
    L.alink_set_train_state.restype = i32
    L.alink_set_train_state.argtypes = [vp, i32, i32, i32, C.POINTER(vp)]
    L.alink_roc_counts.restype = i32
    L.alink_roc_counts.argtypes = [vp, vp, vp, i64, i64, i64, i32, vp, i32, i32, vp, vp, C.POINTER(u64), C.POINTER(u64)]
    L.alink_watchdog_record.restype = i32
    L.alink_watchdog_record.argtypes = [vp, C.POINTER(u64)]
    L.alink_disparity_hist.restype = i32
    L.alink_disparity_hist.argtypes = [vp, vp, vp, i64, i64, u64, i32, i32, vp]
    L.alink_disparity_select_sharded.restype = i32
    L.alink_disparity_select_sharded.argtypes = [vp, vp, vp, i64, i32, i64, vp, dbl, i32, vp, i32, vp, i32,
                                                 C.POINTER(i64), C.POINTER(i64)]
    L.alink_set_reserved_sms.restype = i32
    L.alink_set_reserved_sms.argtypes = [vp, i32]
    L.alink_refine_detail.restype = i32
    L.alink_refine_detail.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(dbl), C.POINTER(C.c_float), C.POINTER(C.c_int)]
    L.alink_set_refine_eps2.restype = i32
    L.alink_set_refine_eps2.argtypes = [vp, dbl]
    L.alink_refine_report_get.restype = i32
    L.alink_refine_report_get.argtypes = [vp, C.POINTER(RefineReport)]
    L.alink_comm_unique_id.restype = i32
    L.alink_comm_unique_id.argtypes = [vp]
    L.alink_comm_init.restype = i32
    L.alink_comm_init.argtypes = [vp, vp, i32, i32]
    L.alink_comm_destroy.restype = i32
    L.alink_comm_destroy.argtypes = [vp]
    L.alink_comm_push_state.restype = i32
    L.alink_comm_push_state.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.alink_query_sharded.restype = i32
    L.alink_query_sharded.argtypes = [vp, i32, i32, i32, vp, vp, i64, i32, i32, i64, i64, vp, vp, vp, i32, vp]
    L.alink_disparity_select_comm.restype = i32
    L.alink_disparity_select_comm.argtypes = [vp, vp, vp, i64, i32, i64, i64, dbl, dbl, i32, vp, i32, vp, i32,
                                              C.POINTER(i64), C.POINTER(i64)]
    L.alink_get_table_rows.restype = i32
    L.alink_get_table_rows.argtypes = [vp, i32, i64, i64, vp, i32]
    L.alink_perturb_rows.restype = i32
    L.alink_perturb_rows.argtypes = [vp, i32, vp, i64, i32, i32, i32, u64, i64, i32, dbl, i32, i32]
    L.alink_score_column.restype = i32
    L.alink_score_column.argtypes = [vp, i32, i32, vp, i32]
    L.alink_min_disparity.restype = i32
    L.alink_min_disparity.argtypes = [vp, vp, vp, i64, i32, i32, vp, i32]
    L.alink_de_init.restype = i32
    L.alink_de_init.argtypes = [vp, i32, i32, i32, vp, i32, i32, vp, vp, vp, vp, vp]
    L.alink_de_mutate.restype = i32
    L.alink_de_mutate.argtypes = [vp, vp, vp, vp, vp, u64, u64, dbl, dbl, dbl, vp]
    L.alink_de_repair.restype = i32
    L.alink_de_repair.argtypes = [vp, vp, vp]
    L.alink_de_evaluate.restype = i32
    L.alink_de_evaluate.argtypes = [vp, i32, i32, vp, vp]
    L.alink_de_select.restype = i32
    L.alink_de_select.argtypes = [vp, vp]
    L.alink_de_get.restype = i32
    L.alink_de_get.argtypes = [vp, vp, vp]
    _lib = L
    return L


def _is_torch(a):
    return type(a).__module__.startswith("torch")


def _ptr_space(a):
    """(void* address, space) of a numpy array (host) or torch tensor (host or cuda). None -> (None, HOST)."""
    if a is None:
        return None, HOST
    if _is_torch(a):
        assert a.is_contiguous(), "tensor must be contiguous"
        return C.c_void_p(a.data_ptr()), (DEVICE if a.is_cuda else HOST)
    assert a.flags["C_CONTIGUOUS"], "array must be C-contiguous"
    return a.ctypes.data_as(C.c_void_p), HOST


def _space_of(*arrs):
    spaces = {_ptr_space(a)[1] for a in arrs if a is not None}
    if len(spaces) > 1:
        raise AlinkError("all buffers of one call must live in the same memory space")
    return spaces.pop() if spaces else HOST


def _pp(arrs):
    holder = (C.c_void_p * len(arrs))(*[a.ctypes.data for a in arrs])
    return C.cast(holder, C.POINTER(C.c_void_p)), holder


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


FLAG_CHECK_INDICES = 1  # include/alink_b200.h ALINK_FLAG_CHECK_INDICES


class Context:
    """One GPU, one stream, all scratch. Mirrors `alink_ctx`."""

    def __init__(self, device: int = 0, check_indices: bool = False):
        self._L = load()
        h = C.c_void_p()
        rc = self._L.alink_create(C.byref(h), int(device), FLAG_CHECK_INDICES if check_indices else 0)
        if rc != 0:
            raise AlinkError(f"alink_create(device={device}) failed with {rc}: no usable CUDA device "
                             "(this library has no CPU fallback)")
        self._h = h
        self.device = device
        self._keep = {}

    def close(self):
        if getattr(self, "_h", None):
            self._L.alink_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def watchdog_record(self):
        """None, or the records timed-out device waits left behind (list of dicts; see include/alink_b200.h)."""
        r = (C.c_uint64 * 256)()
        self._L.alink_watchdog_record(self._h, r)
        if not r[0]:
            return None
        out = []
        for i in range(min(int(r[0]), 63)):
            w, b = int(r[1 + 2 * i]), int(r[2 + 2 * i])
            out.append({"block": w >> 32, "warp": (w & 0xFFFFFFFF) >> 5, "lane": w & 31, "barrier_smem_addr": b >> 32,
                        "parity": (b >> 8) & 1, "cluster_rank": b & 0xFF})
        prog = [int(r[128 + i]) for i in range(64)]
        if any(prog):
            out.append({"progress(block x role: 0 producer,1 L1,2 L2,4-7 epilogue,8-15 gather)": [prog[16 * b:16 * b + 16] for b in range(4)]})
        return out

    def _check(self, rc):
        if rc != 0:
            msg = f"[{rc}] {self._L.alink_last_error(self._h).decode()}"
            if rc == E_CUDA:
                wd = self.watchdog_record()
                if wd:
                    msg += f" | device watchdog: a wait timed out: {wd}"
            raise AlinkError(msg)

    # ---- state
    def set_heads(self, group, heads, feat_mode=FEAT_ABSDIFF, compute=FP32_EXACT):
        """heads: object with K, D_in, H1, H2, C, act, W1..b3 lists (alink_b200.synth.Heads)."""
        keep, ptrs = [], []
        for lst in (heads.W1, heads.b1, heads.W2, heads.b2, heads.W3, heads.b3):
            arrs = [_f32(a) for a in lst]
            p, holder = _pp(arrs)
            keep.append((arrs, holder))
            ptrs.append(p)
        self._check(self._L.alink_set_heads(self._h, group, heads.K, heads.D_in, heads.H1, heads.H2, heads.C,
                                            ACT_ID[heads.act], feat_mode, *ptrs, compute))

    def set_table(self, slot, emb):
        p, space = _ptr_space(emb)
        if _is_torch(emb):
            import torch
            dtype = {torch.float32: F32, torch.bfloat16: BF16}[emb.dtype]
        else:
            if emb.dtype != np.float32:
                raise AlinkError("host tables must be float32")
            dtype = F32
        M, D = emb.shape
        self._check(self._L.alink_set_table(self._h, slot, p, M, D, dtype, space))
        if space == DEVICE:
            self._keep[("table", slot)] = emb  # borrowed: keep alive

    def set_table_async(self, slot, emb):
        """Upload a HOST table (numpy or pinned CPU torch tensor, float32) on the copy stream and return at once; the
        next call that reads the slot waits for it on the device. The buffer is kept alive until the slot is set again."""
        p, space = _ptr_space(emb)
        if space != HOST:
            raise AlinkError("set_table_async takes a host table")
        if _is_torch(emb):
            import torch
            dtype = {torch.float32: F32, torch.bfloat16: BF16}[emb.dtype]
        else:
            if emb.dtype != np.float32:
                raise AlinkError("host tables must be float32")
            dtype = F32
        M, D = emb.shape
        self._check(self._L.alink_set_table_async(self._h, slot, p, M, D, dtype))
        self._keep[("table", slot)] = emb

    # ---- scoring
    def _outs(self, N, K, Cc, want_per, want_cons, out_device):
        if out_device:
            import torch
            dev = torch.device("cuda", self.device)
            per = torch.empty((N, K, Cc), dtype=torch.float32, device=dev) if want_per else None
            cons = torch.empty((N, Cc), dtype=torch.float32, device=dev) if want_cons else None
        else:
            per = np.empty((N, K, Cc), np.float32) if want_per else None
            cons = np.empty((N, Cc), np.float32) if want_cons else None
        return per, cons

    def score_indexed(self, group, slot_a, slot_b, ia, ib, N, K, Cc, want_per=True, want_cons=True,
                      out_device=False):
        per, cons = self._outs(N, K, Cc, want_per, want_cons, out_device)
        pia, sp = _ptr_space(ia)
        pib, _ = _ptr_space(ib)
        _space_of(ia, ib)
        self._check(self._L.alink_score_indexed(self._h, group, slot_a, slot_b, pia, pib, N, sp,
                                                _ptr_space(per)[0], _ptr_space(cons)[0],
                                                DEVICE if out_device else HOST))
        return per, cons

    def score_dense(self, group, XL, XR, K, Cc, want_per=True, want_cons=True, out_device=False):
        N, D = XL.shape
        per, cons = self._outs(N, K, Cc, want_per, want_cons, out_device)
        pl, sp = _ptr_space(XL)
        pr, _ = _ptr_space(XR)
        _space_of(XL, XR)
        self._check(self._L.alink_score_dense(self._h, group, pl, pr, N, D, sp, _ptr_space(per)[0],
                                              _ptr_space(cons)[0], DEVICE if out_device else HOST))
        return per, cons

    def utility(self, measure, N, want=True):
        util = np.empty((N,), np.float32) if want else None
        self._check(self._L.alink_utility(self._h, MEASURE_ID[measure], _ptr_space(util)[0], HOST))
        return util

    def utility_from_proba(self, per_learner, measure, want_consensus=False):
        per_learner = np.ascontiguousarray(per_learner, dtype=np.float32) if not _is_torch(per_learner) else per_learner
        N, K, Cc = per_learner.shape
        p, sp = _ptr_space(per_learner)
        util = np.empty((N,), np.float32)
        cons = np.empty((N, Cc), np.float32) if want_consensus else None
        self._check(self._L.alink_utility_from_proba(self._h, p, N, K, Cc, MEASURE_ID[measure], sp,
                                                     _ptr_space(util)[0], _ptr_space(cons)[0], HOST))
        return (util, cons) if want_consensus else util

    # ---- selection
    def select_topk(self, B, base=0):
        idx = np.empty((B,), np.int64)
        util = np.empty((B,), np.float32)
        keys = np.empty((B,), np.uint64)
        self._check(self._L.alink_select_topk(self._h, B, base, _ptr_space(idx)[0], _ptr_space(util)[0],
                                              _ptr_space(keys)[0], HOST))
        return idx, util, keys

    def topk_values(self, values, B, base=0):
        if not _is_torch(values):
            values = _f32(values)
        p, sp = _ptr_space(values)
        N = values.shape[0]
        idx = np.empty((B,), np.int64)
        util = np.empty((B,), np.float32)
        keys = np.empty((B,), np.uint64)
        self._check(self._L.alink_topk_values(self._h, p, N, base, B, sp, _ptr_space(idx)[0],
                                              _ptr_space(util)[0], _ptr_space(keys)[0], HOST))
        return idx, util, keys

    def topk_values_noalloc(self, values, B, idx_out=None, util_out=None, keys_out=None, base=0):
        """Top-B through the static-scratch entry point (B <= 4096). values and the outputs are torch CUDA tensors the
        caller allocated (int64 / float32 / uint64-as-int64 of B entries); nothing is allocated and nothing synchronises."""
        for t in (values, idx_out, util_out, keys_out):
            if t is not None and not (_is_torch(t) and t.is_cuda):
                raise AlinkError("alink_topk_values_noalloc takes device tensors only")
        self._check(self._L.alink_topk_values_noalloc(self._h, _ptr_space(values)[0], values.shape[0], base, B,
                                                      _ptr_space(idx_out)[0], _ptr_space(util_out)[0],
                                                      _ptr_space(keys_out)[0]))

    def topk_keys(self, keys, B):
        if not _is_torch(keys):
            keys = np.ascontiguousarray(keys, dtype=np.uint64)
        p, sp = _ptr_space(keys)
        out = np.empty((B,), np.uint64)
        self._check(self._L.alink_topk_keys(self._h, p, keys.shape[0], B, sp, _ptr_space(out)[0], HOST))
        return out

    def query(self, group, slot_a, slot_b, ia, ib, N, measure, B, base=0, timing=False, out=None):
        """Fused step. ia/ib: numpy (host) or torch cuda int32. Returns (idx, util, keys[, Timing])."""
        pia, sp = _ptr_space(ia)
        pib, _ = _ptr_space(ib)
        _space_of(ia, ib)
        if out is None:
            out = (np.empty((B,), np.int64), np.empty((B,), np.float32), np.empty((B,), np.uint64))
        idx, util, keys = out  # numpy (host) or torch cuda tensors (left on the device; any may be None)
        osp = _space_of(idx, util, keys)
        t = Timing() if timing else None
        self._check(self._L.alink_query(self._h, group, slot_a, slot_b, pia, pib, N, sp, MEASURE_ID[measure], B,
                                        base, _ptr_space(idx)[0], _ptr_space(util)[0], _ptr_space(keys)[0],
                                        osp, C.byref(t) if timing else None))
        return (idx, util, keys, t) if timing else (idx, util, keys)

    def disparity_select(self, p2, p1, ratio=0.25, eps=0.05, blind=False, y=None):
        """p2 (T, N), p1 (N,), y (N,) or None: numpy (host) or torch cuda float32 tensors (used in place, no host hop)."""
        dev_in = _is_torch(p2) and p2.is_cuda
        if not dev_in:
            p2, p1 = _f32(p2), _f32(p1)
        T, N = int(p2.shape[0]), int(p2.shape[1])
        yy = None if y is None else (y if _is_torch(y) else _f32(y).reshape(-1))
        if dev_in and yy is not None and not _is_torch(yy):
            import torch
            yy = torch.from_numpy(yy).to(p2.device)
        sp = _space_of(p2, p1) if yy is None else _space_of(p2, p1, yy)
        idx = np.empty((max(N, 1),), np.int64)
        n_out, active = C.c_int64(0), C.c_int64(0)
        self._check(self._L.alink_disparity_select(self._h, _ptr_space(p2)[0], _ptr_space(p1)[0], N, T,
                                                   float(ratio), float(eps), int(bool(blind)),
                                                   _ptr_space(yy)[0], sp, _ptr_space(idx)[0], HOST,
                                                   C.byref(n_out), C.byref(active)))
        return idx[:n_out.value].copy(), int(active.value)

    def disparity_hist(self, q, p1, base, prefix, pass_):
        """One radix-select histogram pass (2048 uint32 counters) over this shard's keys (|q - p1|, base + j)."""
        N = int(q.shape[0])
        (pq, sq), (pp, sp) = _ptr_space(q), _ptr_space(p1)
        if N and sq != sp:
            raise ValueError("q and p1 must live in the same memory space")
        hist = np.zeros((2048,), np.uint32)
        self._check(self._L.alink_disparity_hist(self._h, pq, pp, N, int(base), int(prefix), int(pass_), sq,
                                                 hist.ctypes.data))
        return hist

    def disparity_select_sharded(self, p2, p1, base, thr_keys, eps=0.05, blind=False, y=None):
        """Shard-local part of the A-LINK rule against GLOBAL per-noise threshold keys; returns
        (global indices ascending, ACTIVE_COUNT of this shard)."""
        T, N = int(p2.shape[0]), int(p2.shape[1])
        thr = np.ascontiguousarray(thr_keys, dtype=np.uint64)
        assert thr.shape == (T,)
        (pp2, s2), (pp1, s1) = _ptr_space(p2), _ptr_space(p1)
        py, sy = _ptr_space(y) if y is not None else (None, s2)
        if N and not (s2 == s1 == sy):
            raise ValueError("p2, p1 and y must live in the same memory space")
        idx = np.empty((max(N, 1),), np.int64)
        n_out, active = C.c_int64(0), C.c_int64(0)
        self._check(self._L.alink_disparity_select_sharded(self._h, pp2, pp1, N, T, int(base), thr.ctypes.data,
                                                           float(eps), int(bool(blind)), py, s2, idx.ctypes.data, HOST,
                                                           C.byref(n_out), C.byref(active)))
        return idx[:n_out.value].copy(), int(active.value)

    def train_step(self, group, learner, slot_a, slot_b, ia, ib, y, lr=1.0, rho=0.95, eps=1e-7, apply=True,
                   want_probs=False, C_out=2):
        """One Adadelta minibatch step (or, apply=False, an evaluation) -> (loss, accuracy[, probs])."""
        nb = int(len(y))
        yy = y if _is_torch(y) else _f32(y).reshape(-1)
        sp = _space_of(ia, ib, yy)
        loss, acc = C.c_float(0), C.c_float(0)
        probs = np.empty((nb, C_out), np.float32) if want_probs else None
        self._check(self._L.alink_train_step(self._h, group, learner, slot_a, slot_b, _ptr_space(ia)[0], _ptr_space(ib)[0],
                                             _ptr_space(yy)[0], nb, sp, float(lr), float(rho), float(eps), int(bool(apply)),
                                             C.byref(loss), C.byref(acc), _ptr_space(probs)[0]))
        return (float(loss.value), float(acc.value), probs) if want_probs else (float(loss.value), float(acc.value))

    def train_epoch(self, group, learner, slot_a, slot_b, ia, ib, y, batch, lr=1.0, rho=0.95, eps=1e-7, apply=True):
        """ceil(n / batch) minibatch steps enqueued back to back, one host round trip -> (loss[nbatch], acc[nbatch]);
        bit-identical to a loop of train_step over the same batches."""
        n = int(len(y))
        yy = y if _is_torch(y) else _f32(y).reshape(-1)
        sp = _space_of(ia, ib, yy)
        nbatch = (n + int(batch) - 1) // int(batch)
        out = np.empty((nbatch, 2), np.float32)
        self._check(self._L.alink_train_epoch(self._h, group, learner, slot_a, slot_b, _ptr_space(ia)[0], _ptr_space(ib)[0],
                                              _ptr_space(yy)[0], n, int(batch), sp, float(lr), float(rho), float(eps),
                                              int(bool(apply)), _ptr_space(out)[0]))
        return out[:, 0].copy(), out[:, 1].copy()

    def get_train_state(self, group, learner, which, shapes):
        """Six arrays (Keras order) of learner `learner`: which = 0 parameters, 1 Adadelta E[g^2], 2 E[dx^2]."""
        arrs = [np.empty(shp, np.float32) for shp in shapes]
        pp, keep = _pp(arrs)
        self._check(self._L.alink_get_train_state(self._h, group, learner, which, pp))
        return arrs

    def set_train_state(self, group, learner, which, arrs):
        arrs = [_f32(a) for a in arrs]
        pp, keep = _pp(arrs)
        self._check(self._L.alink_set_train_state(self._h, group, learner, which, pp))

    def roc_counts(self, scores, mask, thresholds, roc_case, row0=0):
        """(tp[nthr], fp[nthr], n_genuine, n_impostor) over the strict upper triangle of a score-matrix block."""
        P, G = int(scores.shape[0]), int(scores.shape[1])
        thr = np.ascontiguousarray(thresholds, dtype=np.float64).reshape(-1)
        if not _is_torch(scores):
            scores = np.ascontiguousarray(scores, dtype=np.float32)
        if not _is_torch(mask):
            mask = np.ascontiguousarray(mask, dtype=np.int32)
        sp = _space_of(scores, mask)
        tp, fp = np.zeros(len(thr), np.uint64), np.zeros(len(thr), np.uint64)
        ng, ni = C.c_uint64(0), C.c_uint64(0)
        self._check(self._L.alink_roc_counts(self._h, _ptr_space(scores)[0], _ptr_space(mask)[0], P, G, int(row0),
                                             int(roc_case), thr.ctypes.data, len(thr), sp, tp.ctypes.data, fp.ctypes.data,
                                             C.byref(ng), C.byref(ni)))
        return tp, fp, int(ng.value), int(ni.value)

    def score_allpairs(self, group, slot_probe, slot_gallery, row0, nrows, G, K, Cc, want_per=False, want_cons=True):
        """Scores of probe rows [row0, row0+nrows) against all G gallery rows: consensus (nrows*G, C)."""
        per, cons = self._outs(nrows * G, K, Cc, want_per, want_cons, False)
        self._check(self._L.alink_score_allpairs(self._h, group, slot_probe, slot_gallery, row0, nrows,
                                                 _ptr_space(per)[0], _ptr_space(cons)[0], HOST))
        return per, cons

    def query_allpairs(self, group, slot_probe, slot_gallery, row0, nrows, measure, B, timing=False):
        idx, util, keys = np.empty((B,), np.int64), np.empty((B,), np.float32), np.empty((B,), np.uint64)
        t = Timing() if timing else None
        self._check(self._L.alink_query_allpairs(self._h, group, slot_probe, slot_gallery, row0, nrows,
                                                 MEASURE_ID[measure], B, _ptr_space(idx)[0], _ptr_space(util)[0],
                                                 _ptr_space(keys)[0], HOST, C.byref(t) if timing else None))
        return (idx, util, keys, t) if timing else (idx, util, keys)

    def query_allpairs_keys(self, group, slot_probe, slot_gallery, row0, nrows, measure, B):
        return self.query_allpairs(group, slot_probe, slot_gallery, row0, nrows, measure, B)

    def set_refine_eps(self, eps):
        self._check(self._L.alink_set_refine_eps(self._h, float(eps)))

    # ---- multi-GPU exchange behind the C ABI (csrc/comm.cuh)
    @staticmethod
    def comm_unique_id():
        """128 bytes from ncclGetUniqueId (rank 0), to be shipped to every rank."""
        buf = C.create_string_buffer(128)
        rc = load().alink_comm_unique_id(buf)
        if rc != 0:
            raise AlinkError(f"[{rc}] NCCL is not available (libnccl.so.2 could not be loaded)")
        return buf.raw

    def comm_init(self, uid, rank, world):
        self._check(self._L.alink_comm_init(self._h, C.c_char_p(uid), int(rank), int(world)))

    def comm_destroy(self):
        self._check(self._L.alink_comm_destroy(self._h))

    def comm_push_state(self):
        """(enabled, used by the last query_sharded): the peer-push exchange of csrc/comm.cuh."""
        en, used = C.c_int(0), C.c_int(0)
        self._check(self._L.alink_comm_push_state(self._h, C.byref(en), C.byref(used)))
        return bool(en.value), bool(used.value)

    def query_sharded(self, group, slot_a, slot_b, ia, ib, N, measure, B, base=0, timing=False):
        """alink_query on this rank's shard + NCCL all-gather of the keys + merge, all inside the library: every rank
        returns the same GLOBAL (idx, util, keys)."""
        pia, sp = _ptr_space(ia)
        pib, _ = _ptr_space(ib)
        _space_of(ia, ib)
        idx, util, keys = np.empty((B,), np.int64), np.empty((B,), np.float32), np.empty((B,), np.uint64)
        t = Timing() if timing else None
        self._check(self._L.alink_query_sharded(self._h, group, slot_a, slot_b, pia, pib, N, sp, MEASURE_ID[measure], B, base,
                                                idx.ctypes.data, util.ctypes.data, keys.ctypes.data, HOST,
                                                C.byref(t) if timing else None))
        return (idx, util, keys, t) if timing else (idx, util, keys)

    def disparity_select_comm(self, p2, p1, base, n_total, ratio=0.25, eps=0.05, blind=False, y=None):
        """The A-LINK rule on this rank's shard of a pool of n_total pairs, thresholds selected on the device across the
        communicator (collective call). Returns (this shard's selected GLOBAL indices ascending, global ACTIVE_COUNT)."""
        dev_in = _is_torch(p2) and p2.is_cuda
        if not dev_in:
            p2, p1 = _f32(p2), _f32(p1)
        T, N = int(p2.shape[0]), int(p2.shape[1])
        yy = None if y is None else (y if _is_torch(y) else _f32(y).reshape(-1))
        if dev_in and yy is not None and not _is_torch(yy):
            import torch
            yy = torch.from_numpy(yy).to(p2.device)
        sp = _space_of(p2, p1) if yy is None else _space_of(p2, p1, yy)
        idx = np.empty((max(N, 1),), np.int64)
        n_out, active = C.c_int64(0), C.c_int64(0)
        self._check(self._L.alink_disparity_select_comm(self._h, _ptr_space(p2)[0], _ptr_space(p1)[0], N, T, int(base), int(n_total),
                                                        float(ratio), float(eps), int(bool(blind)), _ptr_space(yy)[0], sp,
                                                        idx.ctypes.data, HOST, C.byref(n_out), C.byref(active)))
        return idx[:n_out.value].copy(), int(active.value)

    # ---- the A2-LINK noise-augmented pool on the device (csrc/perturb.cuh)
    def perturb_rows(self, slot_src, idx, N, side, t, seed, slot_dst, pair_base=0, ncoord=40, vmax=0.25, normalize=True):
        p, sp = _ptr_space(idx)
        self._check(self._L.alink_perturb_rows(self._h, slot_src, p, int(N), sp, int(side), int(t), int(seed), int(pair_base),
                                               int(ncoord), float(vmax), int(bool(normalize)), slot_dst))

    def get_table(self, slot, nrows, D, row0=0, out=None):
        """fp32 rows [row0, row0 + nrows) of a table slot -> host array (or a torch cuda tensor `out`)."""
        if out is None:
            out = np.empty((int(nrows), int(D)), np.float32)
        p, sp = _ptr_space(out)
        self._check(self._L.alink_get_table_rows(self._h, int(slot), int(row0), int(nrows), p, sp))
        return out

    def score_column(self, learner, col, out=None, N=None):
        """Column of the last scores (learner -1 = Bagging mean) into a torch cuda float32 tensor `out`, or a new host array."""
        if out is None:
            out = np.empty((N,), np.float32)
        p, sp = _ptr_space(out)
        self._check(self._L.alink_score_column(self._h, int(learner), int(col), p, sp))
        return out

    def min_disparity(self, p2, p1, want=False):
        """util[j] = min_t |p2[t][j] - p1[j]|, left in the context for select_topk (p2 (T, N), p1 (N,): host or cuda)."""
        T, N = int(p2.shape[0]), int(p2.shape[1])
        sp = _space_of(p2, p1)
        u = np.empty((N,), np.float32) if want else None
        self._check(self._L.alink_min_disparity(self._h, _ptr_space(p2)[0], _ptr_space(p1)[0], N, T, sp, _ptr_space(u)[0], HOST))
        return u

    # ---- batched differential evolution (csrc/de.cuh; alink_b200/attack.py drives it)
    def de_init(self, pop, slot_left, slot_right, il, ir, targets, lo, hi):
        pop = np.ascontiguousarray(pop, dtype=np.float64)
        P, n, G = pop.shape
        il, ir, tg = (np.ascontiguousarray(a, dtype=np.int32) for a in (il, ir, targets))
        lo, hi = np.ascontiguousarray(lo, dtype=np.float64), np.ascontiguousarray(hi, dtype=np.float64)
        assert il.shape == ir.shape == tg.shape == (P,) and lo.shape == hi.shape == (G,)
        self._check(self._L.alink_de_init(self._h, P, n, G, pop.ctypes.data, slot_left, slot_right, il.ctypes.data,
                                          ir.ctypes.data, tg.ctypes.data, lo.ctypes.data, hi.ctypes.data))
        self._de_shape = (P, n, G)

    def de_mutate_reference(self, scale, r0, r1, cross=None):
        """Trials from the caller's (the reference's) random stream; returns the per-problem count of out-of-range genes."""
        P, n, G = self._de_shape
        scale = np.ascontiguousarray(scale, dtype=np.float64)
        r0, r1 = np.ascontiguousarray(r0, dtype=np.int32), np.ascontiguousarray(r1, dtype=np.int32)
        assert scale.shape == (P,) and r0.shape == r1.shape == (P, n)
        cr = None if cross is None else np.ascontiguousarray(cross, dtype=np.uint8)
        nv = np.zeros((P,), np.uint32)
        self._check(self._L.alink_de_mutate(self._h, scale.ctypes.data, r0.ctypes.data, r1.ctypes.data,
                                            None if cr is None else cr.ctypes.data, 0, 0, 0.0, 0.0, 1.0, nv.ctypes.data))
        return nv

    def de_mutate_device(self, seed, gen, dither=(0.5, 1.0), recombination=1.0):
        self._check(self._L.alink_de_mutate(self._h, None, None, None, None, int(seed), int(gen), float(dither[0]),
                                            float(dither[1]), float(recombination), None))

    def de_repair(self, rnd, off):
        rnd = np.ascontiguousarray(rnd, dtype=np.float64)
        off = np.ascontiguousarray(off, dtype=np.int64)
        self._check(self._L.alink_de_repair(self._h, rnd.ctypes.data if rnd.size else None, off.ctypes.data))

    def de_evaluate(self, group, which, K=1, Cc=2, D=None, want_rows=False):
        """which: 0 population (first evaluation), 1 trials, 2 best members -> (probs (P, K, C), rows (P, D) or None)."""
        if which != 2:
            self._check(self._L.alink_de_evaluate(self._h, group, which, None, None))
            return None
        P = self._de_shape[0]
        probs = np.empty((P, K, Cc), np.float32)
        rows = np.empty((P, D), np.float32) if want_rows else None
        self._check(self._L.alink_de_evaluate(self._h, group, 2, probs.ctypes.data, None if rows is None else rows.ctypes.data))
        return probs, rows

    def de_select(self, active=None):
        a = None if active is None else np.ascontiguousarray(active, dtype=np.uint8)
        self._check(self._L.alink_de_select(self._h, None if a is None else a.ctypes.data))

    def de_get(self, want_pop=True):
        P, n, G = self._de_shape
        pop = np.empty((P, n, G), np.float64) if want_pop else None
        e = np.empty((P, n), np.float32)
        self._check(self._L.alink_de_get(self._h, None if pop is None else pop.ctypes.data, e.ctypes.data))
        return pop, e

    def set_refine_eps2(self, eps2):
        self._check(self._L.alink_set_refine_eps2(self._h, C.c_double(float(eps2))))

    def refine_report(self):
        """dict of what the last FP32_REFINE query did (include/alink_b200.h: alink_refine_report)."""
        r = RefineReport()
        self._check(self._L.alink_refine_report_get(self._h, C.byref(r)))
        return r.as_dict()

    def refine_stats(self):
        """(candidates re-scored exactly, max |approx - exact| utility among them) of the last query."""
        n, e = C.c_int64(0), C.c_float(0)
        self._check(self._L.alink_refine_stats(self._h, C.byref(n), C.byref(e)))
        return int(n.value), float(e.value)

    def refine_detail(self):
        """dict(filter, eps_used, rms_err, rounds) of the last FP32_REFINE query."""
        f, e, r, n = C.c_int(0), C.c_double(0), C.c_float(0), C.c_int(0)
        self._check(self._L.alink_refine_detail(self._h, C.byref(f), C.byref(e), C.byref(r), C.byref(n)))
        return {"filter": "f16" if f.value else "bf16", "eps_used": float(e.value), "rms_err": float(r.value),
                "rounds": int(n.value)}

    def set_stream(self, cuda_stream):
        """cuda_stream: integer cudaStream_t (torch.cuda.current_stream().cuda_stream) or None."""
        self._check(self._L.alink_set_stream(self._h, C.c_void_p(cuda_stream) if cuda_stream else None))

    def set_reserved_sms(self, n):
        """Leave n SMs free of the persistent head kernels (for kernels of other streams: feed barriers, collectives)."""
        self._check(self._L.alink_set_reserved_sms(self._h, int(n)))

    def sync(self):
        self._check(self._L.alink_sync(self._h))

    @property
    def launches(self):
        return int(self._L.alink_launch_count(self._h))


_default_ctx = {}


def default_context(device: int | None = None) -> Context:
    """Process-wide context per device (the reference's single global TF session, code/ALINK.py:22-25)."""
    if device is None:
        device = int(os.environ.get("LOCAL_RANK", "0")) if os.environ.get("ALINK_USE_LOCAL_RANK", "1") == "1" else 0
    if device not in _default_ctx:
        _default_ctx[device] = Context(device)
    return _default_ctx[device]
