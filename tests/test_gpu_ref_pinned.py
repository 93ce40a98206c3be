"""CUDA path, through the drop-in classes, against fixtures PRODUCED BY THE REFERENCE'S OWN CODE
(tests/golden/ref_*.npz; see tests/test_ref_pinned.py for what is pinned and what is not).
The per-learner probabilities (row S1-S3, the Keras Dense graph) are the CUDA head's, which is bit-exact with the
oracle's in FP32_EXACT (checked here by hash against the fixture's input hash); everything downstream — Bagging mean,
measures, samplers, Committee containers, the A-LINK rule — must then equal what the reference's own lines returned."""
import os

import numpy as np
import pytest

from alink_b200 import alink, committee, engine, siamese, siamese3, uncertainty
from alink_b200.keras_wrapper import KerasClassifier
from alink_b200.learners import ActiveLearner, Committee
from test_ref_pinned import GOLD, bits, case, check_vec, mrf, sha

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _fp32_mode():
    engine.set_compute_mode("fp32")
    yield
    engine.set_compute_mode("fp32")


def _nets(pool, module=siamese):
    nets = []
    for k in range(pool.heads.K):
        net = module.SiameseNetwork((pool.D,), "/tmp/alink_ref_pinned_%d" % k, 0.1)
        net.set_weights(*pool.heads.learner(k))
        nets.append(net)
    return nets


def _dense(pool):
    return [np.ascontiguousarray(pool.table[pool.ia]), np.ascontiguousarray(pool.table[pool.ib])]


@pytest.mark.parametrize("name", ["cfg1", "cfg2_50k"])
def test_cuda_measures_and_samplers_equal_reference(name):
    fx, pool, _ = case(name)
    B = int(fx["B"])
    nets = _nets(pool)
    X = engine.PairPool(pool.table, pool.ia, pool.ib)
    bag = committee.Bagging(nets, [])
    assert sha(bag.vote_proba(X)) == str(fx["input_sha256"])          # CUDA scores == the fixture's inputs, bit for bit
    cons = bag.predict(X)                                             # code/ALINK.py:160 -> committee.py:13-20
    check_vec(fx, "consensus", cons)
    check_vec(fx, "classifier_uncertainty", uncertainty.classifier_uncertainty(bag, X))
    check_vec(fx, "classifier_margin", uncertainty.classifier_margin(bag, X))
    check_vec(fx, "classifier_entropy", uncertainty.classifier_entropy(bag, X), exact=False, tol=2e-7)
    check_vec(fx, "single_uncertainty", uncertainty.classifier_uncertainty(nets[0], X))
    for nm, fn in (("uncertainty", uncertainty.uncertainty_sampling), ("margin", uncertainty.margin_sampling)):
        idx, rows = fn(bag, X, n_instances=B)
        np.testing.assert_array_equal(np.sort(idx), fx["sel_" + nm])
        np.testing.assert_array_equal(rows[0], pool.table[pool.ia[idx]])
        np.testing.assert_array_equal(rows[1], rows[0])               # (sic) left twice, uncertainty.py:159
    idx, _ = uncertainty.entropy_sampling(bag, X, n_instances=B)
    u = uncertainty.classifier_entropy(bag, X)
    edge = np.sort(u)[::-1][B - 1]
    near = {int(i) for i in np.nonzero(np.abs(u - edge) <= 4e-7)[0]}   # own logf vs scipy: <= 2e-7 per value
    assert (set(idx.tolist()) ^ set(fx["sel_entropy"].tolist())) <= near
    idx, _ = uncertainty.uncertainty_sampling(nets[0], X, n_instances=B)
    np.testing.assert_array_equal(np.sort(idx), fx["single_sel_uncertainty"])
    X.close()


def test_refine_mode_selection_equals_reference_sampler():
    """FP32_REFINE (the bench default) on the 50 k slice: the same SET the reference's uncertainty_sampling picked."""
    fx, pool, _ = case("cfg2_50k")
    engine.set_compute_mode("refine")
    bag = committee.Bagging(_nets(pool), [])
    X = engine.PairPool(pool.table, pool.ia, pool.ib)
    for nm, fn in (("uncertainty", uncertainty.uncertainty_sampling), ("margin", uncertainty.margin_sampling)):
        idx, _ = fn(bag, X, n_instances=int(fx["B"]))
        np.testing.assert_array_equal(np.sort(idx), fx["sel_" + nm])
    X.close()


@pytest.mark.parametrize("name,module", [("cfg1", siamese), ("cfg2_50k", siamese), ("cfg1_sigmoid", siamese3)])
def test_cuda_committee_containers_equal_reference(name, module):
    fx, pool, per = case(name)
    X = _dense(pool)
    learners = [ActiveLearner(estimator=KerasClassifier(lambda n=n: n), query_strategy=uncertainty.uncertainty_sampling)
                for n in _nets(pool, module)]
    com = Committee(learners)
    check_vec(fx, "com_vote_proba", com.vote_proba(X))                # learners.py:383-416 (float64 container)
    check_vec(fx, "com_predict_proba", com.predict_proba(X))          # :345 float64 mean
    check_vec(fx, "com_predict", com.predict(X))                      # :316-332
    check_vec(fx, "com_kc_predict_proba", learners[0].estimator.predict_proba(X))  # keras_wrapper.py:263-270
    check_vec(fx, "com_kc_predict", np.asarray(learners[0].estimator.predict(X)))  # :235-242
    if per.shape[2] == 2:
        check_vec(fx, "com_vote", com.vote(X))                        # :365-381


@pytest.mark.parametrize("name", ["cfg1", "cfg2_50k"])
def test_cuda_disparity_rule_equals_alink_py_block(name):
    fx, pool, per = case(name)
    p2, p1, y = mrf.disparity_inputs(pool, per)
    for tag in ("r25", "r50", "blind"):
        ratio, eps, blind = fx[f"disp_{tag}_params"]
        idx, active = alink.disparity_select(np.ascontiguousarray(p2[:, :, 1]), np.ascontiguousarray(p1[:, 1]), float(ratio),
                                             float(eps), bool(blind), y[:, 0])
        np.testing.assert_array_equal(idx, fx[f"disp_{tag}_idx"])
        assert active == int(fx[f"disp_{tag}_active"])
