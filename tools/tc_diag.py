"""Diagnostic run of the head kernel in the debug build: one scoring call with progress counters on; on a failure
prints how long the call took, the device watchdog records and every role's progress."""
import os
import sys
import time

import numpy as np

os.environ["ALINK_LIB_DEBUG"] = "1"
os.environ.setdefault("ALINK_TC_DBG_DELAY", "64:0")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from alink_b200 import _ffi, synth  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
K = int(sys.argv[2]) if len(sys.argv) > 2 else 2
pool = synth.make_pool(1, N=max(N, 64), D=512, K=K, M=800)
ctx = _ffi.Context(0)
ctx.set_table(0, pool.table)
ctx.set_heads(0, pool.heads, compute=_ffi.BF16_TC)
t0 = time.time()
try:
    per, _ = ctx.score_indexed(0, 0, 0, pool.ia[:N].copy(), pool.ib[:N].copy(), N, pool.heads.K, pool.heads.C)
    print("ok in %.3f s" % (time.time() - t0), "mean p", float(np.mean(per)))
except Exception as e:  # noqa: BLE001
    print("FAILED after %.3f s:" % (time.time() - t0), e)
    print("watchdog:", ctx.watchdog_record())
