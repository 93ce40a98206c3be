"""Adversarial pool generator on the GPU (csrc/de.cuh, alink_b200/attack.py) against the REFERENCE SOLVER's own output
(tests/golden/ref_de.npz: /root/reference/code/differential_evolution.py run in the build container, see
tests/golden/make_ref_de_fixture.py): with the reference's random stream the device populations, energies and best
members are identical for 10 generations on 64 pairs (attack.py's settings), and for binomial crossover."""
import hashlib
import os

import numpy as np
import pytest

from alink_b200 import attack, engine, siamese
from test_attack_cpu import GOLD, mk

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _fp32_mode():
    engine.set_compute_mode("fp32")
    yield
    engine.set_compute_mode("fp32")


def _net(pool):
    net = siamese.SiameseNetwork((pool.D,), "/tmp/alink_attack_model", 0.1)
    net.set_weights(*pool.heads.learner(0))
    return net


@pytest.mark.parametrize("scenario", ["A", "B"])
def test_device_de_equals_the_reference_solver(scenario):
    fx = np.load(os.path.join(GOLD, "ref_de.npz"))
    pool, targets = mk.problem_inputs()
    kw = mk.SCENARIOS[scenario]
    probs = kw["problems"]
    net = _net(pool)
    bounds = [(0, pool.D), (-0.25, 0.25)] * kw["coord_count"]
    popmul = max(1, kw["popsize"] // len(bounds))
    de = attack.BatchedDE(engine.specs_of(net), pool.table, pool.table, pool.ia[probs], pool.ib[probs], targets[probs], bounds,
                          popsize=popmul, dither=(0.5, 1.0), recombination=kw["recombination"], seed=kw["seed0"], rng="reference")
    traj = []
    for _ in range(kw["maxiter"]):
        de.step()
        traj.append(de.state()[1][:, 0].copy())
    pop, e = de.state()
    np.testing.assert_array_equal(e.view(np.uint32), fx[f"{scenario}_energies"].view(np.uint32))
    np.testing.assert_array_equal(pop[:, 0, :], fx[f"{scenario}_best"])
    for j in range(len(probs)):
        assert hashlib.sha256(np.ascontiguousarray(pop[j]).tobytes()).hexdigest() == str(fx[f"{scenario}_pop_sha"][j])
    np.testing.assert_array_equal(np.stack(traj, axis=1), fx[f"{scenario}_traj"])
    # the product of the generator: the perturbed left rows of the best members
    p, rows = de.best()
    lim = np.array(bounds, dtype="float").T
    want = np.stack([attack.perturb((0.5 * (lim[0] + lim[1]) + (pop[j, 0] - 0.5) * np.fabs(lim[0] - lim[1]))[None, :],
                                    pool.table[pool.ia[probs[j]]])[0] for j in range(len(probs))])
    np.testing.assert_array_equal(rows, want)
    np.testing.assert_array_equal(1 - p[np.arange(len(probs)), 0, targets[probs]], e[:, 0])


def test_attack_all_device_stream_reaches_the_target_class():
    """The production mode (device random stream, no host round trip per generation) through the reference's call shape
    (code/attack.py:91-103): every attack flips the model to the target class within the budget, touching <= 40 coordinates."""
    pool, targets = mk.problem_inputs()
    net = _net(pool)
    X = [np.ascontiguousarray(pool.table[pool.ia]), np.ascontiguousarray(pool.table[pool.ib])]
    onehot = np.eye(2, dtype=np.float32)[targets]
    att = attack.EmbeddingAttacker(net)
    rows, de = att.attack_all(X, onehot, coord_count=40, maxiter=20, popsize=250, seed=5, rng="device", return_solver=True)
    assert rows.shape == X[0].shape and rows.dtype == np.float32
    assert np.all((rows != X[0]).sum(axis=1) <= 40)
    pred = np.argmax(net.predict([rows, X[1]]), axis=1)
    assert np.all(pred == targets)
    assert not de.active.any() and de.nit <= 20
    rows2 = att.attack_all(X, onehot, coord_count=40, maxiter=20, popsize=250, seed=5, rng="device")
    np.testing.assert_array_equal(rows, rows2)  # the device stream is a pure function of (seed, problem, generation, member, gene)


def test_noise_objects_through_bagging_attack_model_like_alink_py():
    """code/ALINK.py:93,97,160,170: `ensembleNoise = [cls(model=M2, sess, feature_model)]`, `bag = Bagging(ensemble,
    ensembleNoise)`, `noisy = bag.attackModel(batch_x, IMAGERES, m1_labels)`, then M2 scores [noisy[0][jj], noisy[1][jj]].
    The device-backed noise types: 'coordinate' rows equal the numpy restatement of the counter-based stream bit for bit;
    'adversarial' moves every pair to the class its label names."""
    from alink_b200 import committee, noise
    from oracle import ref_numpy as R
    pool, targets = mk.problem_inputs()
    net = _net(pool)
    n = 24
    X = [np.ascontiguousarray(pool.table[pool.ia[:n]]), np.ascontiguousarray(pool.table[pool.ib[:n]])]
    onehot = np.eye(2, dtype=np.float32)[targets[:n]]
    attacks = [noise.get_relevant_noise("coordinate")(model=net, sess=None, feature_model=None),
               noise.AdversarialNoise(net, None, None, coord_count=40, maxiter=20, popsize=250, seed=5)]
    attacks[0].seed, attacks[0].t = 46, 3
    bag = committee.Bagging([net], attacks)
    noisy = bag.attackModel(X, 224, onehot)
    assert len(noisy[0]) == 2 and len(noisy[1]) == 2
    for side in (0, 1):
        want = R.perturb_rows(X[side], np.arange(n), side, 3, 46)
        np.testing.assert_array_equal(noisy[side][0].view(np.uint32), want.view(np.uint32))
    np.testing.assert_array_equal(noisy[1][1], X[1])  # the embedding attacker perturbs the left side only
    assert np.all((noisy[0][1] != X[0]).sum(axis=1) <= 40)
    pred = np.argmax(net.predict([noisy[0][1], noisy[1][1]]), axis=1)
    np.testing.assert_array_equal(pred, targets[:n])
