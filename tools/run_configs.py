"""Measured runs of the BASELINE.json configs that are not the bench line (they are parity-test shapes; these
timings are evidence for DESIGN.md, not bench values). One GPU = one of the 8 shards of each config.
  python tools/run_configs.py 5     DFW-style all-pairs 50k x 50k x 512-d, probe rows [0, 6250) (1/8 of the matrix)
  python tools/run_configs.py 4     A2-LINK pool: 1 M base pairs x T=16 perturbed tables, M1 = K=5, M2 = K=1, disparity rule
  python tools/run_configs.py k1    single-learner scoring (the HBM-bound case of SURVEY 8d): GB/s of the gather
Prints one JSON line per run."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from alink_b200 import _ffi, synth  # noqa: E402

PEAK_TF, PEAK_GBS = 1409.9, 6535.4
try:
    with open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")) as f:
        _p = json.load(f)
        PEAK_TF, PEAK_GBS = _p.get("bf16_tflops_sustained", PEAK_TF), _p.get("hbm_gbs", PEAK_GBS)
except Exception:
    pass


def flops_per_pair(K, D, H1=512, H2=64, C=2):
    return 2.0 * K * (D * H1 + H1 * H2 + H2 * C)


def timed(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    best = None
    for _ in range(reps):
        t0 = time.perf_counter()
        out = fn()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return best, out


def config5(ctx):
    P = G = 50_000
    rows = P // 8
    pool = synth.make_pool(2, N=4096, M=P)
    ctx.set_table(0, pool.table)
    ctx.set_heads(0, pool.heads, compute=_ffi.BF16_TC)
    dt, (idx, util, keys, t) = timed(lambda: ctx.query_allpairs(0, 0, 0, 0, rows, "uncertainty", 1000, timing=True))
    n = rows * G
    print(json.dumps({"config": "configs[4] DFW all-pairs 50k x 50k x 512-d, K=5, uncertainty top-1000; one of 8 probe-row shards",
                      "pairs": n, "seconds": dt, "pairs_per_s": n / dt, "ms_score": t.ms_score,
                      "tflops": flops_per_pair(5, 512) * n / (t.ms_score / 1e3) / 1e12,
                      "frac_of_sustained_bf16_peak": flops_per_pair(5, 512) * n / (t.ms_score / 1e3) / 1e12 / PEAK_TF,
                      "projected_8gpu_seconds_full_matrix": dt, "top1_pair": [int(idx[0]) // G, int(idx[0]) % G]}))


def k1(ctx, dtype="f32"):
    pool = synth.make_pool(2, N=1_000_000, K=1)
    tab = torch.from_numpy(pool.table).cuda()
    if dtype == "bf16":
        tab = tab.to(torch.bfloat16)
    ctx.set_table(0, tab)
    ctx.set_heads(0, pool.heads, compute=_ffi.BF16_TC)
    ia, ib = torch.from_numpy(pool.ia).cuda(), torch.from_numpy(pool.ib).cuda()
    dt, (_, _, _, t) = timed(lambda: ctx.query(0, 0, 0, ia, ib, pool.N, "uncertainty", 1000, timing=True), reps=5)
    s = 4 if dtype == "f32" else 2
    byts = pool.N * (2 * 512 * s + 8 + 4 * 2 + 4)
    print(json.dumps({"config": f"single learner (K=1), 1 M pairs x 512-d {dtype} table: the HBM-bound case (SURVEY 8d)",
                      "ms_score": t.ms_score, "pairs_per_s": pool.N / (t.ms_score / 1e3),
                      "algorithmic_GBps": byts / (t.ms_score / 1e3) / 1e9, "frac_of_hbm_peak": byts / (t.ms_score / 1e3) / 1e9 / PEAK_GBS,
                      "tflops": flops_per_pair(1, 512) * pool.N / (t.ms_score / 1e3) / 1e12}))


def config4(ctx):
    T = 16
    pool = synth.make_pool(2, N=1_000_000)
    dev = torch.device("cuda", 0)
    base = torch.from_numpy(pool.table).to(dev)
    ia, ib = torch.from_numpy(pool.ia).to(dev), torch.from_numpy(pool.ib).to(dev)
    N, M, D = pool.N, base.shape[0], base.shape[1]
    m2 = synth.make_pool(2, N=4096, K=1).heads  # the disguised-faces model M2: one learner
    ctx.set_heads(0, pool.heads, compute=_ffi.BF16_TC)  # M1 committee, K = 5
    ctx.set_heads(1, m2, compute=_ffi.BF16_TC)
    gen = torch.Generator(device=dev).manual_seed(42 + 4)
    tabs = []
    for t in range(T):  # 40 perturbed coordinates per row (the embedding-space analogue of 40 attacked pixels)
        delta = torch.zeros((M, D), device=dev)
        cols = torch.randint(0, D, (M, 40), device=dev, generator=gen)
        delta.scatter_(1, cols, (torch.rand((M, 40), device=dev, generator=gen) - 0.5) * 0.5)
        tabs.append(torch.nn.functional.normalize(base + delta, dim=1).contiguous())
    p2 = torch.empty((T, N), device=dev, dtype=torch.float32)

    def step():
        ctx.set_table(0, base)
        _, cons = ctx.score_indexed(0, 0, 0, ia, ib, N, 5, 2, want_per=False, want_cons=True, out_device=True)
        p1 = cons[:, 1].contiguous()
        for t in range(T):
            ctx.set_table(1, tabs[t])
            per, _ = ctx.score_indexed(1, 1, 1, ia, ib, N, 1, 2, want_per=True, want_cons=False, out_device=True)
            p2[t] = per.reshape(N, 2)[:, 1]
        return p1, p2

    dt_score, (p1, p2_) = timed(step, reps=2)
    p1h, p2h = p1.cpu().numpy(), p2_.cpu().numpy()
    t0 = time.perf_counter()
    idx, active = ctx.disparity_select(p2h, p1h, 0.25, 0.05, False, None)
    dt_rule = time.perf_counter() - t0
    scored = N + T * N
    print(json.dumps({"config": "configs[3] A2-LINK pool: 1 M base pairs x T=16 perturbed tables; M1 K=5 on the base pairs, M2 K=1 on all 16 M perturbed pairs; disparity rule ratio 0.25 eps 0.05",
                      "pair_scorings": scored, "seconds_scoring": dt_score, "pair_scorings_per_s": scored / dt_score,
                      "seconds_rule_host_buffers": dt_rule, "selected": int(len(idx)), "active_count": int(active)}))


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "5"
    ctx = _ffi.Context(0)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    if which == "5":
        config5(ctx)
    elif which == "4":
        config4(ctx)
    else:
        k1(ctx, "f32")
        k1(ctx, "bf16")
    ctx.close()
