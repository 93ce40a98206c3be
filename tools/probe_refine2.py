"""FP32_REFINE v2 (three stages) against FP32_EXACT on several pools / measures: selection identity, report, timing.
  python tools/probe_refine2.py [N]"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from alink_b200 import _ffi, synth  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
ctx = _ffi.Context(0)
ctx.set_stream(torch.cuda.current_stream().cuda_stream)
dev = torch.device("cuda", 0)
for act, K, gain in (("softmax", 5, 1.0), ("softmax", 3, 4.0), ("sigmoid", 5, 1.0), ("softmax", 5, 16.0)):
    pool = synth.make_pool(2, N=N, K=K, act=act)
    if gain != 1.0:
        for k in range(K):
            pool.heads.W3[k] = (pool.heads.W3[k] * np.float32(gain)).astype(np.float32)
    tab = torch.from_numpy(pool.table).to(dev)
    ia, ib = torch.from_numpy(pool.ia).to(dev), torch.from_numpy(pool.ib).to(dev)
    ctx.set_table(0, tab)
    ctx.set_heads(0, pool.heads, compute=_ffi.FP32_EXACT)
    ctx.set_heads(1, pool.heads, compute=_ffi.FP32_REFINE)
    for measure in ("uncertainty", "neg_margin", "entropy", "consensus_entropy", "kl_max"):
        ie, ue, ke = ctx.query(0, 0, 0, ia, ib, N, measure, 1000)
        best = None
        for rep in range(3):
            ir, ur, kr, t = ctx.query(1, 0, 0, ia, ib, N, measure, 1000, timing=True)
            best = t.ms_total if best is None else min(best, t.ms_total)
        rep_ = ctx.refine_report()
        print(json.dumps({"act": act, "K": K, "gain": gain, "measure": measure, "identical": bool(np.array_equal(ke, kr)),
                          "ms": round(best, 3), "stage_ms": {"filter": round(t.ms_score, 3), "refine": round(t.ms_utility, 3),
                                                              "select": round(t.ms_select, 3)},
                          "launches": t.launches, **{k: (round(v, 9) if isinstance(v, float) else v) for k, v in rep_.items()}}))
ctx.close()
