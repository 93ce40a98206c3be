"""BASELINE configs[3] — the A2-LINK noise-augmented pool — at 1/50 scale (20,000 base pairs x T = 16 perturbed copies)
through the product's device-resident path (alink.a2link_pool_select: alink_perturb_rows -> M2 scores -> alink_score_column
-> alink_disparity_select / alink_min_disparity, no host hop) against the host oracle with the SAME counter-based stream
(oracle/ref_numpy.py perturb_rows + the C oracle head + the C oracle's rule): perturbed rows, p1, p2, the rule's selection
(code/ALINK.py:171-201) and the top-B by min-disparity are bit-identical in FP32_EXACT."""
import numpy as np
import pytest
import torch

from alink_b200 import _ffi, alink, synth
from oracle import ref_numpy as R

pytestmark = pytest.mark.gpu


def test_perturbed_rows_equal_the_numpy_restatement(ctx):
    pool = synth.make_pool(2, N=3000, M=5000)
    ctx.set_table(0, pool.table)
    for side, t, idx in ((0, 0, pool.ia), (1, 7, pool.ib), (0, 15, None)):
        n = pool.N if idx is not None else 3000
        ctx.perturb_rows(0, idx, n, side, t, 46, 3, pair_base=11)
        # read the generated table back through an identity K = 1 gather is overkill: score it against itself instead
        want = R.perturb_rows(pool.table, idx if idx is not None else np.arange(n), side, t, 46, pair_base=11)
        assert np.abs(np.linalg.norm(want, axis=1) - 1).max() < 1e-6
        heads = synth.make_pool(2, N=64, K=1).heads
        ctx.set_heads(3, heads, compute=_ffi.FP32_EXACT)
        base_rows = np.ascontiguousarray(pool.table[:n])
        ctx.set_table(4, base_rows)
        per, _ = ctx.score_indexed(3, 3, 4, None, None, n, 1, 2, want_per=True, want_cons=False)
        from oracle import oracle_ffi
        ref = oracle_ffi.score(want, base_rows, None, None, heads, want_util=False, want_consensus=False)
        np.testing.assert_array_equal(per.view(np.uint32), ref["per_learner"][:n].view(np.uint32))


def test_config4_pool_at_one_fiftieth_scale(ctx, oracle):
    N, T, seed = 20_000, 16, 46
    pool = synth.make_pool(2, N=N, M=8000)
    m2 = synth.make_pool(2, N=4096, K=1, M=8000).heads
    ctx.set_table(0, pool.table)
    ctx.set_heads(0, pool.heads, compute=_ffi.FP32_EXACT)
    ctx.set_heads(1, m2, compute=_ffi.FP32_EXACT)
    ia, ib = torch.from_numpy(pool.ia).cuda(), torch.from_numpy(pool.ib).cuda()
    y = pool.genuine.astype(np.float32)
    idx, active, top_idx, top_util, (p1, p2) = alink.a2link_pool_select(ctx, 0, 1, 0, ia, ib, N, pool.heads.K, T=T, seed=seed, y=y, B=1000)
    # host oracle, same stream
    r1 = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads, want_util=False)
    p1_ref = np.ascontiguousarray(r1["consensus"][:, 1])
    p2_ref = np.empty((T, N), np.float32)
    for t in range(T):
        L = R.perturb_rows(pool.table, pool.ia, 0, t, seed)
        Rr = R.perturb_rows(pool.table, pool.ib, 1, t, seed)
        p2_ref[t] = oracle.score(L, Rr, None, None, m2, want_util=False, want_consensus=False)["per_learner"][:, 0, 1]
    np.testing.assert_array_equal(p1.cpu().numpy().view(np.uint32), p1_ref.view(np.uint32))
    np.testing.assert_array_equal(p2.cpu().numpy().view(np.uint32), p2_ref.view(np.uint32))
    ridx, ractive = oracle.disparity_select(p2_ref, p1_ref, 0.25, 0.05, False, y)
    np.testing.assert_array_equal(idx, ridx)
    assert active == ractive
    u = np.min(np.abs(p2_ref - p1_ref[None, :]), axis=0)
    tidx, tutil, _ = oracle.topk(u, 1000)
    np.testing.assert_array_equal(top_idx, tidx)
    np.testing.assert_array_equal(top_util.view(np.uint32), tutil.view(np.uint32))
    print(f"config 4 at 1/50: rule selects {len(idx)} of {N}, ACTIVE_COUNT {active}; top-1 min-disparity {top_util[0]:.4f}")
