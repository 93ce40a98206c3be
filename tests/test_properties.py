"""Size-independent properties of the path, checked with hypothesis on the CPU oracle: they are the
same properties the GPU tests use at BASELINE's full sizes, where the oracle itself is too slow to be the checker."""
import numpy as np
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

from alink_b200 import synth
from alink_b200.dist import decode_keys, shard_range

_POOL = {}


def _pool(act):
    if act not in _POOL:
        _POOL[act] = synth.make_pool(1, N=600, act=act)
    return _POOL[act]


_S = dict(max_examples=20, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture])


@settings(**_S)
@given(act=st.sampled_from(["softmax", "sigmoid"]), seed=st.integers(0, 2**31 - 1))
def test_probability_rows_and_permutation_equivariance(oracle, act, seed):
    pool = _pool(act)
    out = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads, measure="entropy")
    per = out["per_learner"]
    if act == "softmax":
        assert np.abs(per.sum(axis=2) - 1.0).max() <= 2e-7  # rows sum to 1 (code/siamese.py:32)
    assert (per >= 0).all() and (per <= 1).all()
    perm = np.random.default_rng(seed).permutation(pool.N)
    out2 = oracle.score(pool.table, pool.table, pool.ia[perm].copy(), pool.ib[perm].copy(), pool.heads, measure="entropy")
    np.testing.assert_array_equal(out2["per_learner"], per[perm])  # pairs are independent: bit-exact under reordering
    np.testing.assert_array_equal(out2["util"], out["util"][perm])


@settings(**_S)
@given(p=st.lists(st.floats(0.0, 1.0, width=32), min_size=2, max_size=200))
def test_single_learner_measures_are_monotone_in_distance_from_half(oracle, p):
    """C = 2: uncertainty, -margin and entropy all rank pairs by |p - 0.5| (code/uncertainty.py:82,107,130)."""
    p1 = np.asarray(p, np.float32)
    per = np.stack([1 - p1, p1], axis=1)[:, None, :].astype(np.float32)
    dist = np.abs(p1.astype(np.float64) - 0.5)
    order = np.argsort(dist, kind="stable")
    for m in ("uncertainty", "neg_margin", "entropy"):
        u, _ = oracle.utility(per, m)
        assert (np.diff(u[order].astype(np.float64)) <= 1e-6).all(), m  # farther from 0.5 -> never more useful


@settings(**_S)
@given(u=st.lists(st.floats(-1.0, 1.0, width=32), min_size=1, max_size=300), frac=st.floats(0.0, 1.0),
       shards=st.integers(1, 4))
def test_topk_is_a_sorted_subset_and_shards_merge_to_the_single_shot(oracle, u, frac, shards):
    util = np.asarray(u, np.float32)
    util[::7] = util[0]  # ties
    N = len(util)
    B = max(1, int(frac * N))
    idx, val, keys = oracle.topk(util, B)
    assert len(set(idx.tolist())) == B and idx.min() >= 0 and idx.max() < N         # top-B is a subset of the pool
    assert (np.diff(keys.astype(np.float64)) <= 0).all() and (np.diff(val) <= 0).all()   # sorted, utility descending
    rest = np.setdiff1d(np.arange(N), idx)
    if len(rest):
        assert util[rest].max() <= val.min()
        tied = rest[util[rest] == val.min()]
        if len(tied):  # ties broken by lowest index (north_star)
            assert tied.min() > idx[val == val.min()].max()
    parts = []
    for r in range(shards):
        lo, hi = shard_range(N, r, shards)
        if hi > lo:
            parts.append(oracle.topk(util[lo:hi], min(B, hi - lo), base=lo)[2])
    merged = oracle.topk_keys(np.concatenate(parts), B)
    np.testing.assert_array_equal(merged, keys)
    di, dv = decode_keys(merged)
    np.testing.assert_array_equal(di, idx)


@settings(**_S)
@given(seed=st.integers(0, 2**31 - 1), T=st.integers(1, 4), ratio=st.floats(0.0, 1.0), blind=st.booleans())
def test_disparity_rule_properties(oracle, seed, T, ratio, blind):
    rng = np.random.default_rng(seed)
    N = 300
    p1 = rng.random(N).astype(np.float32)
    p2 = np.clip(p1[None] + 0.3 * rng.standard_normal((T, N)), 0, 1).astype(np.float32)
    y = (rng.random(N) < 0.5).astype(np.float32)
    idx, active = oracle.disparity_select(p2, p1, ratio, 0.05, blind, y)
    idx_noy, active_noy = oracle.disparity_select(p2, p1, ratio, 0.05, blind, None)
    assert (np.diff(idx) > 0).all() and set(idx.tolist()) <= set(idx_noy.tolist())   # the oracle filter only removes
    assert active == active_noy == len(idx_noy)                                     # every grey-zone survivor is a query
    assert all(p1[j] <= 0.45 + 1e-7 or p1[j] >= 0.55 - 1e-7 for j in idx_noy)      # grey zone (code/ALINK.py:195)
    if not blind:
        assert len(idx_noy) <= int(N * ratio)                                       # intersection of T prefixes of that size
        one, _ = oracle.disparity_select(p2[:1], p1, ratio, 0.05, False, None)
        assert set(idx_noy.tolist()) <= set(one.tolist())                           # more noise types only intersect further
