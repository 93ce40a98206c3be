"""Oracle and host logic pinned to fixtures PRODUCED BY THE REFERENCE'S OWN CODE (tests/golden/ref_*.npz, written by
tests/golden/make_ref_fixtures.py, which imports the unmodified /root/reference modules under a modAL / keras stub).
Rows pinned: B1 (committee.py:13-20), U1-U3 (uncertainty.py:15-130), Q1 samplers (:133-217), C1-C4
(learners.py:316-416), W1-W2 (keras_wrapper.py:235-270), P1 (readDFW.py:222-244, readMTP.py:123-135), A1-A2
(ALINK.py:171-201). Bit-exact everywhere except scipy.stats.entropy (<= 2e-7: the oracle's own logf).
Not pinned by the reference (not executable here): the Keras Dense graph S1-S3, whose outputs are these tests' INPUTS."""
import hashlib
import importlib.util
import os

import numpy as np
import pytest

from alink_b200 import pairs, synth
from oracle import ref_numpy as R

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
_spec = importlib.util.spec_from_file_location("make_ref_fixtures", os.path.join(GOLD, "make_ref_fixtures.py"))
mrf = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(mrf)  # imports nothing from /root/reference at module level


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def bits(a):
    a = np.ascontiguousarray(a)
    return a.view({4: np.uint32, 8: np.uint64}[a.dtype.itemsize])


_cache = {}


def case(name):
    """(fixture, pool, oracle per-learner probabilities) — regenerated inputs must hash to what the fixture saw."""
    if name not in _cache:
        fx = np.load(os.path.join(GOLD, f"ref_{name}.npz"))
        pool, per = mrf.case_inputs(name)
        assert sha(per) == str(fx["input_sha256"]), "the oracle's scores on the seeded pool changed: regenerate the fixtures"
        _cache[name] = (fx, pool, per)
    return _cache[name]


def check_vec(fx, key, got, exact=True, tol=0.0):
    """Compare with a fixture vector stored in full (small case) or as sha256 + every 97th element (large case)."""
    if key in fx.files:
        want = fx[key]
        assert got.dtype == want.dtype and got.shape == want.shape
        if exact:
            np.testing.assert_array_equal(bits(got), bits(want))
        else:
            assert np.abs(got.astype(np.float64) - want.astype(np.float64)).max() <= tol
    else:
        assert exact
        np.testing.assert_array_equal(bits(got[::97]), bits(fx[key + "_every97"]))
        assert sha(got) == str(fx[key + "_sha256"])


@pytest.mark.parametrize("name", ["cfg1", "cfg2_50k"])
def test_oracle_measures_equal_reference(oracle, name):
    fx, pool, per = case(name)
    if "per_learner" in fx.files:
        np.testing.assert_array_equal(bits(per), bits(fx["per_learner"]))
    u1, cons = oracle.utility(per, "uncertainty")
    check_vec(fx, "consensus", cons)                                  # Bagging.predict, committee.py:18
    check_vec(fx, "proba_uncertainty", u1)                            # uncertainty.py:26
    check_vec(fx, "classifier_uncertainty", u1)                       # :82
    nm, _ = oracle.utility(per, "neg_margin")
    check_vec(fx, "proba_margin", -nm)                                # :43-44 (the sampler ranks -margin, :183)
    check_vec(fx, "classifier_margin", -nm)
    ent, _ = oracle.utility(per, "entropy")
    check_vec(fx, "proba_entropy", ent, exact=False, tol=2e-7)        # scipy.stats.entropy, :60
    check_vec(fx, "classifier_entropy", ent, exact=False, tol=2e-7)
    u0, _ = oracle.utility(np.ascontiguousarray(per[:, :1, :]), "uncertainty")
    check_vec(fx, "single_uncertainty", u0)


@pytest.mark.parametrize("name", ["cfg1", "cfg2_50k"])
def test_oracle_selections_equal_reference_samplers(oracle, name):
    """uncertainty_sampling / margin_sampling / entropy_sampling (uncertainty.py:133-217) -> multi_argmax: equal as
    SETS (argpartition leaves the order unspecified); the fixtures record that the B-th value is not tied."""
    fx, pool, per = case(name)
    B = int(fx["B"])
    for nm, measure in (("uncertainty", "uncertainty"), ("margin", "neg_margin"), ("entropy", "entropy")):
        assert bool(fx["tie_free_" + nm])
        u, _ = oracle.utility(per, measure)
        idx, _, _ = oracle.topk(u, B)
        if nm == "entropy":  # the oracle's entropy differs from scipy's by <= 2e-7: compare away from the boundary
            want = set(fx["sel_entropy"].tolist())
            got = set(idx.tolist())
            edge = np.sort(u)[::-1][B - 1]
            near = {int(i) for i in np.nonzero(np.abs(u - edge) <= 4e-7)[0]}
            assert (got ^ want) <= near
        else:
            np.testing.assert_array_equal(np.sort(idx), fx["sel_" + nm])
    u0, _ = oracle.utility(np.ascontiguousarray(per[:, :1, :]), "uncertainty")
    idx, _, _ = oracle.topk(u0, B)
    np.testing.assert_array_equal(np.sort(idx), fx["single_sel_uncertainty"])


@pytest.mark.parametrize("name", ["cfg1", "cfg2_50k", "cfg1_sigmoid"])
def test_committee_containers_equal_reference(name):
    """Committee.vote_proba / predict_proba / vote / predict and KerasClassifier.predict_proba / predict on the
    numpy restatement (the product's classes are checked against the same fixtures on the GPU)."""
    fx, pool, per = case(name)
    wide = per if per.shape[2] == 2 else np.concatenate([1 - per, per], axis=2)  # keras_wrapper.py:268
    vp = wide.astype(np.float64)
    check_vec(fx, "com_vote_proba", vp)
    check_vec(fx, "com_predict_proba", np.mean(vp, axis=1))           # learners.py:345, float64 mean
    check_vec(fx, "com_predict", np.arange(2)[np.argmax(np.mean(vp, axis=1), axis=1)])
    check_vec(fx, "com_kc_predict_proba", np.ascontiguousarray(wide[:, 0, :]))
    if per.shape[2] == 2:
        check_vec(fx, "com_vote", np.argmax(wide, axis=2).astype(np.float64))
        check_vec(fx, "com_kc_predict", np.arange(2)[np.argmax(per[:, 0, :], axis=-1)])
    else:
        check_vec(fx, "com_kc_predict", np.arange(2)[(per[:, 0, :] > 0.5).astype("int32")])


def test_pair_enumeration_equals_reference_loops():
    fx = np.load(os.path.join(GOLD, "ref_pairs.npz"))
    for tag in ("a", "b", "c"):
        ps, ds = fx[f"dfw_{tag}_plain"], fx[f"dfw_{tag}_dig"]
        ia, ib, Y = pairs.pair_indices_dfw(ps, ds)   # ids in the fixture = row numbers of [plain..., dig...]
        np.testing.assert_array_equal(ia, fx[f"dfw_{tag}_left"])
        np.testing.assert_array_equal(ib, fx[f"dfw_{tag}_right"])
        np.testing.assert_array_equal(Y[:, 0].astype(np.int64), fx[f"dfw_{tag}_y"])
        ia, ib, Y = pairs.pair_indices_mtp(ps)
        np.testing.assert_array_equal(ia, fx[f"mtp_{tag}_left"])
        np.testing.assert_array_equal(ib, fx[f"mtp_{tag}_right"])
        np.testing.assert_array_equal(Y[:, 0].astype(np.int64), fx[f"mtp_{tag}_y"])
        # the loop restatement the other tests use (oracle/ref_numpy.py) agrees with the reference's loops too
        nid = 0
        plain, dig = [], []
        for n in ps:
            plain.append(np.arange(nid, nid + int(n), dtype=np.float32).reshape(int(n), 1))
            nid += int(n)
        for n in ds:
            dig.append(np.arange(nid, nid + int(n), dtype=np.float32).reshape(int(n), 1))
            nid += int(n)
        (l2, r2), y2 = R.create_mini_batch_dfw(plain, dig)
        np.testing.assert_array_equal(l2[:, 0].astype(np.int64), fx[f"dfw_{tag}_left"])
        np.testing.assert_array_equal(r2[:, 0].astype(np.int64), fx[f"dfw_{tag}_right"])
        (l2, r2), y2 = R.create_mini_batch_mtp(plain)
        np.testing.assert_array_equal(l2[:, 0].astype(np.int64), fx[f"mtp_{tag}_left"])
        np.testing.assert_array_equal(y2[:, 0].astype(np.int64), fx[f"mtp_{tag}_y"])


@pytest.mark.parametrize("name", ["cfg1", "cfg2_50k"])
def test_oracle_disparity_rule_equals_alink_py_block(oracle, name):
    """exec of code/ALINK.py:171-201 vs the oracle's alink_oracle_disparity_select on column 1."""
    fx, pool, per = case(name)
    p2, p1, y = mrf.disparity_inputs(pool, per)
    assert mrf.sha(p2) + mrf.sha(p1) + mrf.sha(y) == str(fx["disp_input_sha256"])
    for tag in ("r25", "r50", "blind"):
        ratio, eps, blind = fx[f"disp_{tag}_params"]
        idx, active = oracle.disparity_select(np.ascontiguousarray(p2[:, :, 1]), np.ascontiguousarray(p1[:, 1]), float(ratio),
                                              float(eps), bool(blind), y[:, 0])
        np.testing.assert_array_equal(idx, fx[f"disp_{tag}_idx"])
        assert active == int(fx[f"disp_{tag}_active"])
