"""Edge cases and error behaviour of the entry points added in round 2 (include/alink_b200.h): bad arguments come back as
ALINK_E_* (an AlinkError with the library's text), never as a device fault; host / device variants agree; without a
communicator alink_query_sharded is alink_query."""
import numpy as np
import pytest
import torch

from alink_b200 import _ffi, synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def small(ctx):
    pool = synth.make_pool(1, N=3000)
    ctx.set_table(0, pool.table)
    ctx.set_heads(0, pool.heads, compute=_ffi.FP32_EXACT)
    return pool


def test_query_sharded_without_a_communicator_is_query(ctx, small):
    a = ctx.query(0, 0, 0, small.ia, small.ib, small.N, "uncertainty", 100, base=7)
    b = ctx.query_sharded(0, 0, 0, small.ia, small.ib, small.N, "uncertainty", 100, base=7)
    for x, y in zip(a, b):
        np.testing.assert_array_equal(x, y)
    with pytest.raises(_ffi.AlinkError):
        ctx.disparity_select_comm(np.zeros((1, 8), np.float32), np.zeros(8, np.float32), 0, 8)  # alink_comm_init first


def test_score_column_and_min_disparity_host_and_device(ctx, small):
    per, cons = ctx.score_indexed(0, 0, 0, small.ia, small.ib, small.N, small.heads.K, 2)
    np.testing.assert_array_equal(ctx.score_column(-1, 1, N=small.N), cons[:, 1])   # Bagging mean, column 1 (ALINK.py:176)
    np.testing.assert_array_equal(ctx.score_column(2, 0, N=small.N), per[:, 2, 0])
    d = torch.empty(small.N, dtype=torch.float32, device="cuda")
    ctx.score_column(0, 1, out=d)
    np.testing.assert_array_equal(d.cpu().numpy(), per[:, 0, 1])
    with pytest.raises(_ffi.AlinkError):
        ctx.score_column(small.heads.K, 1, N=small.N)
    rng = np.random.default_rng(0)
    p2, p1 = rng.random((3, 500)).astype(np.float32), rng.random(500).astype(np.float32)
    u = ctx.min_disparity(p2, p1, want=True)
    np.testing.assert_array_equal(u, np.min(np.abs(p2 - p1[None, :]), axis=0))
    idx, util, _ = ctx.select_topk(10)
    want = np.lexsort((np.arange(500), -u.astype(np.float64)))[:10]
    np.testing.assert_array_equal(idx, want)
    u2 = ctx.min_disparity(torch.from_numpy(p2).cuda(), torch.from_numpy(p1).cuda(), want=True)
    np.testing.assert_array_equal(u2, u)


def test_perturb_rows_argument_checks_and_determinism(ctx, small):
    ctx.perturb_rows(0, small.ia, small.N, 0, 3, 9, 3)
    ctx.set_heads(3, synth.make_pool(1, N=64, K=1).heads, compute=_ffi.FP32_EXACT)
    a, _ = ctx.score_indexed(3, 3, 0, None, None, min(small.N, small.table.shape[0]), 1, 2)
    ctx.perturb_rows(0, torch.from_numpy(small.ia).cuda(), small.N, 0, 3, 9, 4)      # device indices, another slot: same rows
    b, _ = ctx.score_indexed(3, 4, 0, None, None, min(small.N, small.table.shape[0]), 1, 2)
    np.testing.assert_array_equal(a, b)
    ctx.perturb_rows(0, small.ia, small.N, 0, 4, 9, 4)                               # another noise type: other rows
    c, _ = ctx.score_indexed(3, 4, 0, None, None, min(small.N, small.table.shape[0]), 1, 2)
    assert np.abs(a - c).max() > 1e-4
    for bad in (dict(side=2), dict(t=-1), dict(slot_dst=0), dict(ncoord=-1)):
        kw = dict(side=0, t=0, slot_dst=3, ncoord=40)
        kw.update(bad)
        with pytest.raises(_ffi.AlinkError):
            ctx.perturb_rows(0, small.ia, small.N, kw["side"], kw["t"], 9, kw["slot_dst"], ncoord=kw["ncoord"])


def test_de_argument_checks(ctx, small):
    P, n, G = 2, 8, 4
    pop = np.random.default_rng(0).random((P, n, G))
    lo, hi = np.zeros(G), np.ones(G)
    with pytest.raises(_ffi.AlinkError):
        ctx.de_init(pop[:, :3], 0, 0, [0, 1], [0, 1], [0, 1], lo, hi)        # fewer than 5 members (differential_evolution.py:415)
    with pytest.raises(_ffi.AlinkError):
        ctx.de_init(pop[:, :, :3], 0, 0, [0, 1], [0, 1], [0, 1], lo[:3], hi[:3])  # genes come in (index, value) pairs
    ctx.de_init(pop, 0, 0, [0, 1], [5, 6], [1, 0], np.array([0, -0.25] * 2), np.array([512, 0.25] * 2))
    with pytest.raises(_ffi.AlinkError):
        ctx.de_select()                                                       # nothing evaluated yet
    ctx.de_evaluate(0, 0)
    ctx.de_mutate_device(1, 1)
    ctx.de_evaluate(0, 1)
    ctx.de_select()
    pop2, e = ctx.de_get()
    assert pop2.shape == (P, n, G) and np.all((pop2 >= 0) & (pop2 <= 1)) and np.all(np.isfinite(e))
    assert np.all(e[:, 0] <= e.min(axis=1) + 0)                               # slot 0 holds the best member


def test_refine_report_for_a_query_that_does_not_refine(ctx, small):
    ctx.query(0, 0, 0, small.ia, small.ib, small.N, "uncertainty", 10)
    rep = ctx.refine_report()
    assert rep["version"] == 0 and rep["n_exact"] == 0
