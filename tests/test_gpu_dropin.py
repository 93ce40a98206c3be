"""GPU tests of the drop-in Python interfaces, written the way the reference's drivers call them
(/root/reference/code/ALINK.py:96-97,160,170; code/existing_al.py:95-110; code/learners.py:287-416;
utilities/generateMatrixDFW.py:19-33) and checked against the CPU oracle."""
import os

import numpy as np
import pytest

from alink_b200 import alink, committee, disagreement, engine, siamese, siamese3, synth, uncertainty
from alink_b200.keras_wrapper import KerasClassifier
from alink_b200.learners import ActiveLearner, Committee

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _fp32_mode():
    engine.set_compute_mode("fp32")
    yield
    engine.set_compute_mode("fp32")


def _nets(pool, module=siamese):
    nets = []
    for k in range(pool.heads.K):
        net = module.SiameseNetwork((pool.D,), "/tmp/alink_test_model_%d" % k, 0.1)
        net.set_weights(*pool.heads.learner(k))
        nets.append(net)
    return nets


def _dense(pool):
    return [np.ascontiguousarray(pool.table[pool.ia]), np.ascontiguousarray(pool.table[pool.ib])]


def test_siamese_predict_and_bagging(oracle):
    pool = synth.make_pool(1, N=1500)
    X = _dense(pool)
    ref = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads)
    nets = _nets(pool)
    p0 = nets[0].predict(X)  # code/siamese.py:130-131
    assert p0.shape == (pool.N, 2) and p0.dtype == np.float32
    np.testing.assert_array_equal(p0, ref["per_learner"][:, 0, :])
    bag = committee.Bagging(nets, [])  # code/ALINK.py:97
    ens = bag.predict(X)  # code/ALINK.py:160
    np.testing.assert_array_equal(ens, ref["consensus"])


def test_weights_roundtrip_h5(tmp_path):
    pool = synth.make_pool(1, N=64)
    net = siamese.SiameseNetwork((pool.D,), str(tmp_path / "m"), 0.1)
    assert net.maybeLoadFromMemory() is False  # the reference swallows the failure (code/siamese.py:114-119)
    net.set_weights(*pool.heads.learner(0))
    net.save()  # <modelName>.h5, a Keras weight file (code/siamese.py:121-125)
    assert os.path.exists(str(tmp_path / "m") + ".h5")
    net2 = siamese.SiameseNetwork((pool.D,), str(tmp_path / "m"), 0.1)
    assert net2.maybeLoadFromMemory() is True
    X = _dense(pool)
    np.testing.assert_array_equal(net.predict(X), net2.predict(X))


@pytest.mark.parametrize("strategy,measure", [("uncertainty_sampling", "uncertainty"), ("margin_sampling", "neg_margin"),
                                              ("entropy_sampling", "entropy")])
def test_active_learner_query_like_existing_al(oracle, strategy, measure):
    """code/existing_al.py:95-110: KerasClassifier around a sigmoid siamese3 net, learner.query([XL, XR], ...)."""
    pool = synth.make_pool(1, N=1200, K=1, act="sigmoid")
    net = _nets(pool, siamese3)[0]
    wrapped = KerasClassifier(lambda: net)
    learner = ActiveLearner(estimator=wrapped, query_strategy=getattr(uncertainty, strategy), verbose=0)
    X = _dense(pool)
    n = int(len(X[0]) * 0.1)
    query_idx, query_inst = learner.query(X, n_instances=n, verbose=0)
    ref = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads, measure=measure)
    ridx, _, _ = oracle.topk(ref["util"], n)
    np.testing.assert_array_equal(query_idx, ridx)
    np.testing.assert_array_equal(query_inst[0], X[0][query_idx])
    np.testing.assert_array_equal(query_inst[1], X[0][query_idx])  # (sic) left twice, code/uncertainty.py:159
    proba = wrapped.predict_proba(X)
    assert proba.shape == (pool.N, 2)
    np.testing.assert_array_equal(proba[:, 1], ref["per_learner"][:, 0, 0])
    np.testing.assert_array_equal(proba[:, 0], 1 - ref["per_learner"][:, 0, 0])  # code/keras_wrapper.py:268


def test_committee_interfaces(oracle):
    pool = synth.make_pool(1, N=900)
    X = _dense(pool)
    learners = [ActiveLearner(estimator=KerasClassifier(lambda n=n: n), query_strategy=uncertainty.uncertainty_sampling)
                for n in _nets(pool)]
    com = Committee(learners)  # default strategy: vote_entropy_sampling (code/learners.py:287)
    ref = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads, measure="vote_entropy")
    vp = com.vote_proba(X)
    assert vp.shape == (pool.N, pool.heads.K, 2) and vp.dtype == np.float64  # np.zeros default dtype, :398
    np.testing.assert_array_equal(vp.astype(np.float32), ref["per_learner"])
    pp = com.predict_proba(X)  # code/learners.py:345: np.mean of the float64 container (not the fp32 Bagging mean)
    assert pp.dtype == np.float64
    np.testing.assert_array_equal(pp, np.mean(ref["per_learner"].astype(np.float64), axis=1))
    assert np.abs(pp - ref["consensus"]).max() <= 1.2e-7  # one fp32 ulp
    votes = com.vote(X)
    assert votes.shape == (pool.N, pool.heads.K)
    np.testing.assert_array_equal(votes, (ref["per_learner"][:, :, 1] > ref["per_learner"][:, :, 0]).astype(np.float64))
    np.testing.assert_array_equal(com.predict(X), np.argmax(pp, axis=1))
    idx, inst = com.query(X, n_instances=50)
    ridx, _, _ = oracle.topk(ref["util"], 50)
    np.testing.assert_array_equal(idx, ridx)
    for fn, m in ((disagreement.consensus_entropy, "consensus_entropy"), (disagreement.KL_max_disagreement, "kl_max")):
        r = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads, measure=m)
        np.testing.assert_array_equal(fn(com, X), r["util"])


def test_pairpool_query_equals_dense(oracle):
    pool = synth.make_pool(1)
    nets = _nets(pool)
    bag = committee.Bagging(nets, [])
    pp = engine.PairPool(pool.table, pool.ia, pool.ib)
    np.testing.assert_array_equal(bag.predict(pp), bag.predict(_dense(pool)))
    learner = ActiveLearner(estimator=bag, query_strategy=uncertainty.uncertainty_sampling)
    idx, inst = learner.query(pp, n_instances=pool.B)
    ref = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads, measure="uncertainty")
    ridx, _, _ = oracle.topk(ref["util"], pool.B)
    np.testing.assert_array_equal(idx, ridx)
    np.testing.assert_array_equal(inst[0], pool.table[pool.ia[idx]])


def test_alink_select_like_alink_py(oracle):
    """code/ALINK.py:160-201: committee on clean features, M2 on T noisy copies, disparity rule on column 1."""
    pool = synth.make_pool(1, N=800)
    rng = np.random.default_rng(3)
    nets = _nets(pool)
    bag = committee.Bagging(nets[:2], [])
    m2 = nets[2]
    X = _dense(pool)
    noisy = [[X[0] + rng.normal(0, 0.02, X[0].shape).astype(np.float32), X[1]] for _ in range(3)]
    y = (rng.random(pool.N) < 0.5).astype(np.float32)
    idx, active, ens = alink.alink_select(bag, m2, X, noisy, col=1, disparity_ratio=0.25, eps=0.05, y=y)
    p1 = ens[:, 1]
    p2 = np.stack([m2.predict(n)[:, 1] for n in noisy])
    ridx, ractive = oracle.disparity_select(p2, p1, 0.25, 0.05, False, y)
    np.testing.assert_array_equal(idx, ridx)
    assert active == ractive


def test_bf16_mode_through_dropin(oracle):
    pool = synth.make_pool(1, N=1000)
    engine.set_compute_mode("bf16")
    bag = committee.Bagging(_nets(pool), [])
    ens = bag.predict(_dense(pool))
    ref = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads)
    assert np.abs(ens - ref["consensus"]).max() <= 2e-2  # north_star bf16 tolerance


def test_all_pairs_rows_like_generate_matrix(oracle):
    """utilities/generateMatrixDFW.py:28-33: row i of the score matrix = predict([features[i]] * G, features)[:, 0]."""
    pool = synth.make_pool(1, N=64, K=1)
    feats = pool.table[:96]
    net = _nets(pool)[0]
    for i in (0, 17):
        XL = np.ascontiguousarray(np.repeat(feats[i:i + 1], len(feats), axis=0))
        row = net.predict([XL, feats])[:, 0]
        ia = np.full(len(feats), i, np.int32)
        ib = np.arange(len(feats), dtype=np.int32)
        ref = oracle.score(feats, feats, ia, ib, pool.heads)
        np.testing.assert_array_equal(row, ref["per_learner"][:, 0, 0])


def test_committee_default_query_in_refine_mode_equals_fp32():
    """Committee(learner_list).query — vote_entropy_sampling by default (code/learners.py:287) — over a 200 k-pair pool:
    'refine' mode (tensor-core vote filter, csrc/refine.cuh) must return the very same rows as 'fp32'."""
    from alink_b200 import _ffi
    pool = synth.make_pool(2, N=200_000)
    pp = engine.PairPool(pool.table, pool.ia, pool.ib)
    got = {}
    for mode in ("fp32", "refine"):
        engine.set_compute_mode(mode)
        learners = [ActiveLearner(estimator=KerasClassifier(lambda n=n: n), query_strategy=uncertainty.uncertainty_sampling)
                    for n in _nets(pool)]
        com = Committee(learners)
        got[mode], _ = com.query(pp, n_instances=500)
        if mode == "refine":
            rep = pp.ctx.refine_report()
            assert rep["version"] == 2 and rep["fallback"] == 0, rep
    np.testing.assert_array_equal(got["refine"], got["fp32"])
    pp.close()
