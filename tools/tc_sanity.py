"""One bf16 tensor-core scoring call at a given size; prints max |dp| vs the oracle. Separate processes per case
(a launch failure poisons the CUDA context). Usage: python tools/tc_sanity.py N [D]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from alink_b200 import _ffi, synth  # noqa: E402
from oracle import oracle_ffi  # noqa: E402

N = int(sys.argv[1])
D = int(sys.argv[2]) if len(sys.argv) > 2 else 512
pool = synth.make_pool(1, N=max(N, 64), D=D, K=2, M=800)
ctx = _ffi.Context(0)
ctx.set_table(0, pool.table)
ctx.set_heads(0, pool.heads, compute=_ffi.BF16_TC)
per, _ = ctx.score_indexed(0, 0, 0, pool.ia[:N].copy(), pool.ib[:N].copy(), N, pool.heads.K, pool.heads.C)
ref = oracle_ffi.score(pool.table, pool.table, pool.ia[:N].copy(), pool.ib[:N].copy(), pool.heads)
print("N", N, "D", D, "NCTA", os.environ.get("ALINK_TC_NCTA", "2"), "max|dp|", float(np.abs(per - ref["per_learner"]).max()))
