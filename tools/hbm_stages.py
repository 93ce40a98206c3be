"""Drive every HBM-bound stage of the path once, as separate launches, for the ncu pass of profiles/r2_ncu_hbm_stages.*
(north_star: gather / entropy / top-k stages are judged on achieved HBM GB/s from ncu captures).
  python tools/hbm_stages.py            # plain run (must exit 0 before the ncu run of the same command)
Stages, all on configs[1]'s shape (1 M pairs, 512-d fp32 table, K = 5, B = 1000) unless noted:
  1. BF16_TC score (probabilities [N][K][C] kept) -> utility_kernel<5> (entropy)  -> radix select + compaction + sort
  2. FP32_REFINE query (mark / compose / maxerr / small top-k kernels)
  3. K = 1 BF16_TC query on fp32 rows: the HBM-bound head launch of SURVEY 8d (the gather itself)
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from alink_b200 import _ffi, synth  # noqa: E402

N = int(os.environ.get("HBM_STAGES_N", 1_000_000))
pool = synth.make_pool(2, N=N)
ctx = _ffi.Context(0)
ctx.set_stream(torch.cuda.current_stream().cuda_stream)
dev = torch.device("cuda", 0)
tab = torch.from_numpy(pool.table).to(dev)
ia, ib = torch.from_numpy(pool.ia).to(dev), torch.from_numpy(pool.ib).to(dev)
ctx.set_table(0, tab)
for rep in range(2):
    ctx.set_heads(0, pool.heads, compute=_ffi.BF16_TC)
    ctx.score_indexed(0, 0, 0, ia, ib, N, 5, 2, want_per=False, want_cons=False)
    for m in ("uncertainty", "entropy", "kl_max"):
        u = ctx.utility(m, N, want=False)
    idx, util, keys = ctx.select_topk(1000)
    ctx.set_heads(1, pool.heads, compute=_ffi.FP32_REFINE)
    ctx.query(1, 0, 0, ia, ib, N, "uncertainty", 1000)
    k1 = synth.make_pool(2, N=4096, K=1).heads
    ctx.set_heads(2, k1, compute=_ffi.BF16_TC)
    ctx.query(2, 0, 0, ia, ib, N, "uncertainty", 1000)
torch.cuda.synchronize()
print("ok", int(idx[0]), float(util[0]))
ctx.close()
