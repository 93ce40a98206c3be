"""BASELINE.json configs[1] at its FULL size (1,000,000 pairs, 512-d, M = 200,000, K = 5, top-1000) through
size-independent properties (the oracle needs minutes for the whole pool, so it checks only the selected rows):

  * FP32_REFINE (the bench default) selects exactly what FP32_EXACT selects — keys, indices, utility bits;
  * the utilities of the selected pairs are bit-identical to the CPU oracle run on just those pairs;
  * ragged shards of the pool, merged by key, equal the single-shot selection;
  * permuting the pairs permutes the selection (equivariance; the tie rule is on the ORIGINAL index, so the
    comparison is on the (utility, pair) multiset);
  * the selection is sorted by (utility desc, index asc) with distinct in-range indices.
"""
import numpy as np
import pytest

from alink_b200 import _ffi, synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pool():
    return synth.make_pool(2)


@pytest.fixture(scope="module")
def single_shot(ctx, pool):
    assert pool.N == 1_000_000 and pool.D == 512 and pool.heads.K == 5 and pool.B == 1000
    ctx.set_table(0, pool.table)
    ctx.set_heads(2, pool.heads, compute=_ffi.FP32_REFINE)
    return ctx.query(2, 0, 0, pool.ia, pool.ib, pool.N, "uncertainty", pool.B)


def test_refine_equals_exact_at_full_size(ctx, pool, single_shot):
    idx, util, keys = single_shot
    ctx.set_heads(1, pool.heads, compute=_ffi.FP32_EXACT)
    eidx, eutil, ekeys = ctx.query(1, 0, 0, pool.ia, pool.ib, pool.N, "uncertainty", pool.B)
    np.testing.assert_array_equal(keys, ekeys)
    np.testing.assert_array_equal(idx, eidx)
    np.testing.assert_array_equal(util.view(np.uint32), eutil.view(np.uint32))
    assert np.all(keys[:-1] > keys[1:])
    assert idx.min() >= 0 and idx.max() < pool.N and len(set(idx.tolist())) == pool.B
    assert np.all((util[:-1] > util[1:]) | ((util[:-1] == util[1:]) & (idx[:-1] < idx[1:])))


def test_selected_utilities_match_the_oracle(ctx, oracle, pool, single_shot):
    idx, util, _ = single_shot
    ref = oracle.score(pool.table, pool.table, pool.ia[idx], pool.ib[idx], pool.heads, measure="uncertainty")
    np.testing.assert_array_equal(util.view(np.uint32), ref["util"].view(np.uint32))
    # and a random sample of NON-selected pairs is at or below the B-th utility (oracle-side check of the cut)
    rng = np.random.default_rng(0)
    rest = np.setdiff1d(rng.choice(pool.N, 4000, replace=False), idx)
    r2 = oracle.score(pool.table, pool.table, pool.ia[rest], pool.ib[rest], pool.heads, measure="uncertainty")
    assert float(r2["util"].max()) <= float(util[-1])


def test_ragged_shards_merge_to_the_single_shot(ctx, pool, single_shot):
    _, _, keys = single_shot
    cuts = [0, 1, 333_333, 333_334 + 401_007, pool.N]  # a 1-pair shard and two uneven ones
    parts = []
    for lo, hi in zip(cuts[:-1], cuts[1:]):
        _, _, k = ctx.query(2, 0, 0, pool.ia[lo:hi], pool.ib[lo:hi], hi - lo, "uncertainty", min(pool.B, hi - lo), base=lo)
        parts.append(k)
    merged = ctx.topk_keys(np.concatenate(parts), pool.B)
    np.testing.assert_array_equal(merged, keys)


def test_permutation_equivariance(ctx, pool, single_shot):
    idx, util, _ = single_shot
    perm = np.random.default_rng(1).permutation(pool.N)
    pidx, putil, _ = ctx.query(2, 0, 0, np.ascontiguousarray(pool.ia[perm]), np.ascontiguousarray(pool.ib[perm]),
                               pool.N, "uncertainty", pool.B)
    back = perm[pidx]
    # distinct utilities select the same pairs; only pairs tied AT the cut may differ
    cut = util[-1]
    above = util > cut
    assert set(idx[above].tolist()) == set(back[putil > cut].tolist())
    np.testing.assert_array_equal(np.sort(util.view(np.uint32)), np.sort(putil.view(np.uint32)))
