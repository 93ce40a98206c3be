"""Per-kernel times of the finetune step (csrc/train.cuh): two passes of alink_train_epoch over 208 pairs, batch 16 —
run plain for the wall time and under `ncu --metrics gpu__time_duration.sum` for the launch list."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from alink_b200 import _ffi, synth  # noqa: E402

n, batch = 208, 16
pool = synth.make_pool(1, N=n, M=400, K=1)
y = np.asarray(pool.genuine[:n], dtype=np.float32)
ia, ib = pool.ia[:n].copy(), pool.ib[:n].copy()
ctx = _ffi.Context(0)
ctx.set_table(0, pool.table)
ctx.set_heads(3, pool.heads, compute=_ffi.FP32_EXACT)
for _ in range(3):
    ctx.train_epoch(3, 0, 0, 0, ia, ib, y, batch, lr=0.1)
t0 = time.perf_counter()
reps = int(os.environ.get("TRAIN_PROBE_REPS", 10))
for _ in range(reps):
    ls, accs = ctx.train_epoch(3, 0, 0, 0, ia, ib, y, batch, lr=0.1)
dt = (time.perf_counter() - t0) / reps
print(f"alink_train_epoch: {dt * 1e3:.3f} ms per pass of {len(ls)} batches = {dt * 1e6 / len(ls):.1f} us per batch; loss {ls[0]:.4f} -> {ls[-1]:.4f}")
ctx.close()
