"""GPU parity: the CUDA path, called through the C ABI, against the CPU oracle on the same seeded
inputs. FP32_EXACT mode must be BIT-EXACT (scores, utilities, selected indices)."""
import numpy as np
import pytest

from alink_b200 import _ffi, synth

pytestmark = pytest.mark.gpu

MEASURES = ("uncertainty", "neg_margin", "entropy", "vote_entropy", "consensus_entropy", "kl_max")


def _install(ctx, pool, group=0, feat_mode=0):
    ctx.set_heads(group, pool.heads, feat_mode=feat_mode, compute=_ffi.FP32_EXACT)
    ctx.set_table(0, pool.table)


@pytest.mark.parametrize("act", ["softmax", "sigmoid"])
@pytest.mark.parametrize("N", [1, 127, 2000])
def test_scores_bit_exact(ctx, oracle, act, N):
    pool = synth.make_pool(1, N=max(N, 64), act=act)
    ia, ib = pool.ia[:N].copy(), pool.ib[:N].copy()
    _install(ctx, pool)
    per, cons = ctx.score_indexed(0, 0, 0, ia, ib, N, pool.heads.K, pool.heads.C)
    ref = oracle.score(pool.table, pool.table, ia, ib, pool.heads)
    assert per.shape == ref["per_learner"].shape
    np.testing.assert_array_equal(per.view(np.uint32), ref["per_learner"].view(np.uint32))
    np.testing.assert_array_equal(cons.view(np.uint32), ref["consensus"].view(np.uint32))
    if act == "softmax":
        np.testing.assert_allclose(per.sum(axis=2), 1.0, atol=2e-7)


@pytest.mark.parametrize("measure", MEASURES)
def test_utilities_bit_exact(ctx, oracle, measure):
    pool = synth.make_pool(1)
    _install(ctx, pool)
    ctx.score_indexed(0, 0, 0, pool.ia, pool.ib, pool.N, pool.heads.K, pool.heads.C, False, False)
    util = ctx.utility(measure, pool.N)
    ref = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads, measure=measure)
    np.testing.assert_array_equal(util.view(np.uint32), ref["util"].view(np.uint32))


@pytest.mark.parametrize("measure", MEASURES)
@pytest.mark.parametrize("act", ["softmax", "sigmoid"])
def test_query_selection_identical(ctx, oracle, measure, act):
    pool = synth.make_pool(1, act=act)
    _install(ctx, pool)
    idx, util, keys, t = ctx.query(0, 0, 0, pool.ia, pool.ib, pool.N, measure, pool.B, timing=True)
    ref = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads, measure=measure)
    ridx, rutil, rkeys = oracle.topk(ref["util"], pool.B)
    np.testing.assert_array_equal(idx, ridx)
    np.testing.assert_array_equal(util.view(np.uint32), rutil.view(np.uint32))
    np.testing.assert_array_equal(keys, rkeys)
    assert t.launches > 0 and t.ms_total > 0


def test_dense_equals_indexed(ctx, oracle):
    pool = synth.make_pool(1, N=777)
    _install(ctx, pool)
    XL, XR = pool.table[pool.ia], pool.table[pool.ib]
    per_d, cons_d = ctx.score_dense(0, np.ascontiguousarray(XL), np.ascontiguousarray(XR), pool.heads.K, pool.heads.C)
    per_i, cons_i = ctx.score_indexed(0, 0, 0, pool.ia, pool.ib, pool.N, pool.heads.K, pool.heads.C)
    np.testing.assert_array_equal(per_d, per_i)
    np.testing.assert_array_equal(cons_d, cons_i)


def test_hadamard_feature_mode(ctx, oracle):
    pool = synth.make_pool(1, N=300, feat_mode=1)
    _install(ctx, pool, feat_mode=1)
    per, _ = ctx.score_indexed(0, 0, 0, pool.ia, pool.ib, pool.N, pool.heads.K, pool.heads.C)
    ref = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads, feat_mode=1)
    np.testing.assert_array_equal(per.view(np.uint32), ref["per_learner"].view(np.uint32))


def test_smallres_widths_and_d2048(ctx, oracle):
    pool = synth.make_pool(1, N=300, D=2048, kind="resnet50", H1=128, H2=32, K=2)
    _install(ctx, pool)
    per, _ = ctx.score_indexed(0, 0, 0, pool.ia, pool.ib, pool.N, pool.heads.K, pool.heads.C)
    ref = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads)
    np.testing.assert_array_equal(per.view(np.uint32), ref["per_learner"].view(np.uint32))


@pytest.mark.parametrize("n,B", [(1, 1), (5, 5), (1000, 1), (4097, 4096), (100000, 1000), (300000, 5000),
                                 # the one-CTA paths (topk.cuh SELECT_SMALL_N = 32768, shared-memory sort up to 4096 keys)
                                 # and their boundaries: fused top-B, small select + global sort, multi-CTA select
                                 (2, 1), (3, 2), (10000, 1000), (32768, 4096), (32768, 32768), (20000, 5000),
                                 (32769, 4096), (32769, 1)])
def test_topk_values(ctx, oracle, n, B):
    rng = np.random.default_rng(n + B)
    v = rng.standard_normal(n).astype(np.float32)
    v[rng.integers(0, n, size=n // 3 + 1)] = 0.25  # heavy ties
    if n > 10:
        v[3] = -0.0
        v[4] = 0.0
    idx, util, keys = ctx.topk_values(v, B, base=7)
    ridx, rutil, rkeys = oracle.topk(v, B, base=7)
    np.testing.assert_array_equal(idx, ridx)
    np.testing.assert_array_equal(keys, rkeys)
    np.testing.assert_array_equal(util, rutil)


@pytest.mark.parametrize("n", [50000, 9000])  # multi-CTA passes / the one-CTA kernel
def test_topk_all_tied_picks_lowest_indices(ctx, n):
    v = np.full(n, 0.673012, np.float32)  # vote-entropy style plateau
    idx, _, _ = ctx.topk_values(v, 1000)
    np.testing.assert_array_equal(idx, np.arange(1000))


def test_topk_keys_merge(ctx, oracle):
    rng = np.random.default_rng(5)
    v = rng.random(40000).astype(np.float32)
    parts = []
    for r in range(4):
        lo, hi = r * 10000, (r + 1) * 10000
        _, _, k = ctx.topk_values(v[lo:hi], 500, base=lo)
        parts.append(k)
    merged = ctx.topk_keys(np.concatenate(parts), 500)
    _, _, ref = oracle.topk(v, 500)
    np.testing.assert_array_equal(merged, ref)


@pytest.mark.parametrize("blind", [False, True])
@pytest.mark.parametrize("with_y", [False, True])
def test_disparity_select(ctx, oracle, blind, with_y):
    rng = np.random.default_rng(11)
    N, T = 5000, 4
    p1 = rng.random(N).astype(np.float32)
    p2 = np.clip(p1[None, :] + rng.normal(0, 0.2, size=(T, N)), 0, 1).astype(np.float32)
    p2[:, :50] = p1[:50]  # zero-disparity ties
    y = (rng.random(N) < 0.5).astype(np.float32) if with_y else None
    idx, active = ctx.disparity_select(p2, p1, 0.25, 0.05, blind, y)
    ridx, ractive = oracle.disparity_select(p2, p1, 0.25, 0.05, blind, y)
    np.testing.assert_array_equal(idx, ridx)
    assert active == ractive


def test_errors_are_reported_not_fatal(ctx):
    pool = synth.make_pool(1, N=64)
    _install(ctx, pool)
    with pytest.raises(_ffi.AlinkError):
        ctx.query(0, 0, 0, pool.ia, pool.ib, pool.N, "uncertainty", pool.N + 1)
    fresh = _ffi.Context(0)  # the session context is shared: other test files may have filled its groups
    fresh.set_table(0, pool.table)
    with pytest.raises(_ffi.AlinkError):
        fresh.query(3, 0, 0, pool.ia, pool.ib, pool.N, "uncertainty", 1)  # group 3 never set
    fresh.close()
    # the context is still usable afterwards
    idx, _, _ = ctx.query(0, 0, 0, pool.ia, pool.ib, pool.N, "uncertainty", 8)
    assert len(idx) == 8


def test_async_table_upload_matches_the_blocking_one(oracle):
    """alink_set_table_async: the upload runs on the copy stream, the next reader waits on the device; alternating
    two slots (the e2e pipeline of bench.py) must give the scores of the blocking upload, also when a slot is
    overwritten right after it was read."""
    import torch

    ctx = _ffi.Context(0)
    pool = synth.make_pool(1, N=3000, M=900)
    t0 = torch.from_numpy(pool.table)
    tabs = [t0.clone().pin_memory(), torch.roll(t0, 7, dims=0).contiguous().pin_memory()]
    K, Cc = pool.heads.K, pool.heads.C
    ctx.set_heads(0, pool.heads, compute=_ffi.FP32_EXACT)
    ref = []
    for t in tabs:
        ctx.set_table(0, t)
        ref.append(ctx.score_indexed(0, 0, 0, pool.ia, pool.ib, pool.N, K, Cc)[0].copy())
    assert not np.array_equal(ref[0], ref[1])
    cur = 1
    ctx.set_table(cur, tabs[0])
    for step in range(6):
        nxt = 3 - cur
        ctx.set_table_async(nxt, tabs[(step + 1) % 2])  # in flight while `cur` is scored
        per, _ = ctx.score_indexed(0, cur, cur, pool.ia, pool.ib, pool.N, K, Cc)
        assert np.array_equal(per.view(np.uint32), ref[step % 2].view(np.uint32)), f"step {step}"
        cur = nxt
    ctx.close()


def test_index_range_check_flag(oracle):
    """ALINK_FLAG_CHECK_INDICES: an out-of-range pair index is an ALINK_E_INVALID error, not a device fault;
    in-range pools are unaffected (bit-exact) and the context stays usable after the error."""
    pool = synth.make_pool(1, N=3000)
    c = _ffi.Context(0, check_indices=True)
    try:
        c.set_table(0, pool.table)
        c.set_heads(0, pool.heads, compute=_ffi.FP32_EXACT)
        good = c.query(0, 0, 0, pool.ia, pool.ib, pool.N, "uncertainty", 50)
        for bad_value in (pool.M, -1, 2**31 - 1):
            for side in (0, 1):
                ia, ib = pool.ia.copy(), pool.ib.copy()
                (ia, ib)[side][1234] = bad_value
                with pytest.raises(_ffi.AlinkError, match="holds index"):
                    c.query(0, 0, 0, ia, ib, pool.N, "uncertainty", 50)
                with pytest.raises(_ffi.AlinkError, match="holds index"):
                    c.score_indexed(0, 0, 0, ia, ib, pool.N, pool.heads.K, pool.heads.C)
        again = c.query(0, 0, 0, pool.ia, pool.ib, pool.N, "uncertainty", 50)
        np.testing.assert_array_equal(good[2], again[2])
        ref = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads, measure="uncertainty")
        np.testing.assert_array_equal(good[0], oracle.topk(ref["util"], 50)[0])
    finally:
        c.close()
    from alink_b200 import engine
    bad = pool.ia.copy()
    bad[7] = pool.M
    with pytest.raises(IndexError, match="holds index"):
        engine.PairPool(pool.table, bad, pool.ib)
