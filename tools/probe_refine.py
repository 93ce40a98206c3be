"""Measures the FP32_REFINE filter error (max / rms |approx - exact| utility over the candidates) per
measure and pool shape, for the half and bf16 operand formats. Run on the GPU box:
  python tools/probe_refine.py            (half filter)
  ALINK_REFINE_FMT=1 python tools/probe_refine.py   (bf16 filter)
The default bounds in csrc/alink_b200.cu (REFINE_EPS_F16 / REFINE_EPS_BF16) come from this output."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from alink_b200 import _ffi, synth  # noqa: E402

SCALE = {"uncertainty": 1.0, "neg_margin": 2.0, "entropy": 1.25, "consensus_entropy": 1.25, "kl_max": 5.0}


def main():
    ctx = _ffi.Context(0)
    ctx.set_refine_eps(8e-3)  # wide band: many candidates -> a good sample of the error distribution
    cases = [("cfg2 1M x 512-d K=5 softmax", dict(cfg=2, N=1_000_000)),
             ("cfg2 300k sigmoid K=3", dict(cfg=2, N=300_000, K=3, act="sigmoid")),
             ("resnet50 2048-d 200k K=5", dict(cfg=1, N=200_000, D=2048, kind="resnet50", K=5, M=20_000)),
             ("smallres 128/32 200k K=5", dict(cfg=1, N=200_000, K=5, H1=128, H2=32, M=20_000))]
    for name, kw in cases:
        cfg = kw.pop("cfg")
        pool = synth.make_pool(cfg, **kw)
        ctx.set_table(0, pool.table)
        ctx.set_heads(0, pool.heads, compute=_ffi.FP32_REFINE)
        for measure in SCALE:
            t0 = time.perf_counter()
            ctx.query(0, 0, 0, pool.ia, pool.ib, pool.N, measure, min(1000, pool.N // 4))
            dt = time.perf_counter() - t0
            n, mx = ctx.refine_stats()
            d = ctx.refine_detail()
            print(f"{name:32s} {measure:18s} filter={d['filter']:4s} survivors={n:8d} max_err={mx:.3e} "
                  f"rms={d['rms_err']:.3e} max/scale={mx / SCALE[measure]:.3e} rounds={d['rounds']} {dt * 1e3:.1f} ms",
                  flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
