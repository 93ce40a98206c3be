"""Pin the C oracle (oracle/alink_oracle.c + include/alink_math.h) against an INDEPENDENT float64
numpy/scipy transcription of the reference lines (oracle/ref_numpy.py) and the committed golden
vectors (tests/golden/, produced by tests/golden/make_golden.py from that transcription)."""
import os

import numpy as np
import pytest

from alink_b200 import synth
from oracle import ref_numpy as R

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TOL_P = 1e-5  # north_star: pair scores within 1e-5 in fp32 mode


def _ref_preds(pool, act, feat_mode=0, dtype=np.float64):
    XL, XR = pool.table[pool.ia], pool.table[pool.ib]
    return [R.head_forward(XL, XR, *pool.heads.learner(k), act=act, feat_mode=feat_mode, dtype=dtype)
            for k in range(pool.heads.K)]


@pytest.mark.parametrize("act", ["softmax", "sigmoid"])
@pytest.mark.parametrize("feat_mode", [0, 1])
def test_scores_match_numpy_transcription(oracle, act, feat_mode):
    pool = synth.make_pool(1, N=512, act=act, feat_mode=feat_mode)
    out = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads, feat_mode=feat_mode)
    preds = _ref_preds(pool, act, feat_mode)
    per = np.stack(preds, axis=1)
    assert np.abs(per - out["per_learner"]).max() < TOL_P
    cons = R.bagging_predict(preds)
    assert np.abs(cons - out["consensus"]).max() < TOL_P
    # fp32 numpy (the dtype Keras would use) agrees too
    per32 = np.stack(_ref_preds(pool, act, feat_mode, np.float32), axis=1)
    assert np.abs(per32 - out["per_learner"]).max() < TOL_P


@pytest.mark.parametrize("measure", R.MEASURES)
@pytest.mark.parametrize("act", ["softmax", "sigmoid"])
def test_measures_match_scipy(oracle, measure, act):
    pool = synth.make_pool(1, N=1000, act=act)
    out = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads, measure=measure)
    # measures on the ORACLE's own probabilities isolate the measure arithmetic
    preds = [out["per_learner"][:, k, :].astype(np.float64) for k in range(pool.heads.K)]
    u = R.utility(measure, preds)
    assert np.abs(u - out["util"]).max() < 2e-6
    u2, _ = oracle.utility(out["per_learner"], measure)
    np.testing.assert_array_equal(u2, out["util"])


def test_single_learner_measures(oracle):
    pool = synth.make_pool(1, N=300, K=1)
    for m in ("uncertainty", "neg_margin", "entropy"):
        out = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads, measure=m)
        u = R.utility(m, [out["per_learner"][:, 0, :].astype(np.float64)])
        assert np.abs(u - out["util"]).max() < 2e-6


def test_vote_entropy_value_set(oracle):
    pool = synth.make_pool(1, K=5, N=2000)
    out = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads, measure="vote_entropy")
    vals = np.unique(out["util"])
    expect = np.array([0.0, 0.5004024, 0.6730117])  # H(0/5), H(1/5), H(2/5): SURVEY App. B
    assert len(vals) <= 3 and all(np.abs(expect - v).min() < 1e-6 for v in vals)


def test_math_header_against_libm(oracle):
    xs = np.concatenate([np.linspace(-86, 88, 3001), [0.0, -0.0, 1e-8, -1e-8]]).astype(np.float32)
    e = np.array([oracle.expf(x) for x in xs], dtype=np.float64)
    ref = np.exp(xs.astype(np.float64))
    assert np.max(np.abs(e - ref) / ref) < 2.5e-7
    assert oracle.expf(0.0) == 1.0
    ps = np.exp(np.linspace(np.log(2e-38), np.log(8.0), 3001)).astype(np.float32)
    l = np.array([oracle.logf(p) for p in ps], dtype=np.float64)
    ref = np.log(ps.astype(np.float64))
    assert np.max(np.abs(l - ref) / np.maximum(np.abs(ref), 1e-2)) < 3e-7
    assert oracle.logf(1.0) == 0.0


def test_topk_rule_and_key_packing(oracle):
    rng = np.random.default_rng(3)
    v = rng.standard_normal(5000).astype(np.float32)
    v[rng.integers(0, 5000, 2000)] = 0.5
    v[7], v[8] = 0.0, -0.0
    idx, util, keys = oracle.topk(v, 700, base=11)
    ref = R.topk_lowest_index(v, 700)
    np.testing.assert_array_equal(idx, ref + 11)
    np.testing.assert_array_equal(util, v[ref] + 0.0)
    assert np.all(np.diff(keys.astype(np.float64)) < 0) or np.all(keys[:-1] > keys[1:])
    # set-equality with modAL's multi_argmax whenever the boundary is not tied
    v2 = rng.permutation(5000).astype(np.float32)
    idx2, _, _ = oracle.topk(v2, 123)
    assert set(idx2.tolist()) == set(R.multi_argmax(v2, 123).tolist())
    # shard merge == single shot
    parts = [oracle.topk(v[lo:lo + 1250], 700 if 700 <= 1250 else 1250, base=lo)[2] for lo in range(0, 5000, 1250)]
    merged = oracle.topk_keys(np.concatenate(parts), 700)
    np.testing.assert_array_equal(merged, oracle.topk(v, 700)[2])


@pytest.mark.parametrize("blind", [False, True])
def test_disparity_rule_matches_python_transcription(oracle, blind):
    rng = np.random.default_rng(4)
    N, T = 1200, 3
    p1 = rng.random(N).astype(np.float32)
    p2 = np.clip(p1[None] + rng.normal(0, 0.25, (T, N)), 0, 1).astype(np.float32)
    p2[:, :40] = p1[:40]
    y = (rng.random(N) < 0.5).astype(np.float32)
    for yy in (None, y):
        idx, active = oracle.disparity_select(p2, p1, 0.25, 0.05, blind, yy)
        ridx, ractive = R.disparity_select(p2, p1, 0.25, 0.05, blind, yy)
        np.testing.assert_array_equal(idx, ridx)
        assert active == ractive


def test_edge_cases(oracle):
    pool = synth.make_pool(1, N=64)
    out = oracle.score(pool.table, pool.table, pool.ia[:0], pool.ib[:0], pool.heads)
    assert out["per_learner"].shape == (0, 3, 2)
    out = oracle.score(pool.table, pool.table, pool.ia[:5], pool.ib[:5], pool.heads)  # ragged (< register block)
    full = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads)
    np.testing.assert_array_equal(out["per_learner"], full["per_learner"][:5])
    # identical pair -> |a-b| = 0 -> the head sees only its biases
    same = oracle.score(pool.table, pool.table, np.array([3], np.int32), np.array([3], np.int32), pool.heads)
    assert np.isfinite(same["per_learner"]).all()
    # thread count must not change a single bit
    one = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads, threads=1)
    np.testing.assert_array_equal(one["per_learner"], full["per_learner"])


@pytest.mark.parametrize("name", ["softmax_d512_k3", "sigmoid_d512_k5", "softmax_d2048_k5"])
def test_golden_vectors(oracle, name):
    z = np.load(os.path.join(GOLD, name + ".npz"))
    kw = {k: z[k].item() for k in ("cfg", "N", "K", "D", "act", "kind", "M")}
    pool = synth.make_pool(int(kw["cfg"]), N=int(kw["N"]), K=int(kw["K"]), D=int(kw["D"]), act=str(kw["act"]),
                           kind=str(kw["kind"]), M=int(kw["M"]))
    # the generator itself must not drift
    assert abs(float(pool.table.astype(np.float64).sum()) - float(z["table_sum"])) < 1e-6 * max(1.0, abs(float(z["table_sum"])))
    np.testing.assert_array_equal(pool.ia, z["ia"])
    out = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads)
    assert np.abs(out["per_learner"] - z["per_learner"]).max() < TOL_P
    assert np.abs(out["consensus"] - z["consensus"]).max() < TOL_P
    for m in R.MEASURES:
        u = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads, measure=m)["util"]
        assert np.abs(u - z["util_" + m]).max() < 1e-5, m
        # top-B by the golden utilities and by the oracle's agree wherever the boundary gap is > 2e-5
        g = z["util_" + m]
        B = 16
        order = R.topk_lowest_index(g, len(g))
        if g[order[B - 1]] - g[order[B]] > 2e-5:
            assert set(oracle.topk(u, B)[0].tolist()) == set(order[:B].tolist()), m


@pytest.mark.parametrize("act", ["softmax", "sigmoid"])
def test_scores_match_a_torch_cpu_twin(oracle, act):
    """BASELINE.md §4's torch-CPU twin: the Dense / relu / softmax (sigmoid) graph of code/siamese.py:24-33 (siamese3.py:18-26)
    in torch's own fp32 CPU kernels (MKL GEMM, its softmax / sigmoid) — a third implementation, independent of both the C
    oracle and the numpy transcription. Not the reference (Keras / TF is not installable here), but it bounds how far any
    faithful fp32 execution of the graph can sit from the oracle: within north_star's 1e-5."""
    import torch
    pool = synth.make_pool(1, N=2000, act=act)
    out = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads)
    xl, xr = torch.from_numpy(pool.table[pool.ia]), torch.from_numpy(pool.table[pool.ib])
    for k in range(pool.heads.K):
        W1, b1, W2, b2, W3, b3 = (torch.from_numpy(a) for a in pool.heads.learner(k))
        x = torch.abs(xl - xr)
        h = torch.relu(torch.relu(x @ W1 + b1) @ W2 + b2)
        z = h @ W3 + b3
        p = torch.softmax(z, dim=1) if act == "softmax" else torch.sigmoid(z)
        assert float((p - torch.from_numpy(out["per_learner"][:, k, :])).abs().max()) < TOL_P
