"""Three FP32_REFINE queries on configs[1] (1 M pairs, 512-d, K = 5, top-1000) — the command of the ncu launch list
profiles/r2_launches_step.txt (tools/launch_step.py prints the last full step)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from alink_b200 import _ffi, synth  # noqa: E402

N = int(os.environ.get("REFINE_STEP_N", 1_000_000))
pool = synth.make_pool(2, N=N)
ctx = _ffi.Context(0)
ctx.set_stream(torch.cuda.current_stream().cuda_stream)
dev = torch.device("cuda", 0)
tab = torch.from_numpy(pool.table).to(dev)
ia, ib = torch.from_numpy(pool.ia).to(dev), torch.from_numpy(pool.ib).to(dev)
ctx.set_table(0, tab)
ctx.set_heads(0, pool.heads, compute=_ffi.FP32_REFINE)
for _ in range(4):
    idx, util, keys, t = ctx.query(0, 0, 0, ia, ib, N, os.environ.get("REFINE_STEP_MEASURE", "uncertainty"), 1000, timing=True)
print(t.as_dict(), ctx.refine_report())
ctx.close()
