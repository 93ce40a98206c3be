import sys, numpy as np
sys.path.insert(0, '/root/repo')
from alink_b200 import _ffi, synth
from oracle import oracle_ffi
ctx = _ffi.Context(0)
for N in (1, 127, 128, 2000):
  for act in ("softmax", "sigmoid"):
    pool = synth.make_pool(1, N=max(N, 64), act=act)
    ia, ib = pool.ia[:N].copy(), pool.ib[:N].copy()
    ctx.set_heads(1, pool.heads, compute=_ffi.BF16_TC)
    ctx.set_table(0, pool.table)
    try:
        per, cons = ctx.score_indexed(1, 0, 0, ia, ib, N, pool.heads.K, pool.heads.C)
        ctx.sync()
        ref = oracle_ffi.score(pool.table, pool.table, ia, ib, pool.heads)
        print(N, act, "err", float(np.abs(per - ref["per_learner"]).max()), flush=True)
    except Exception as e:
        print(N, act, "FAIL", e, flush=True); sys.exit(1)
