"""Adversarial pool generator, host side (no GPU): the product's ReferenceStream (alink_b200/attack.py) + the parallel
formulation of one DE generation (oracle/ref_numpy.py de_*, what csrc/de.cuh implements) + the oracle head reproduce the
REFERENCE SOLVER's own output (tests/golden/ref_de.npz, written by tests/golden/make_ref_de_fixture.py from the unmodified
/root/reference/code/differential_evolution.py) bit for bit."""
import hashlib
import importlib.util
import os

import numpy as np
import pytest

from alink_b200 import attack
from oracle import ref_numpy as R

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
_spec = importlib.util.spec_from_file_location("make_ref_de_fixture", os.path.join(GOLD, "make_ref_de_fixture.py"))
mk = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(mk)


def run_numpy(oracle, pool, targets, p, coord_count, popsize, maxiter, recombination, seed0):
    D = pool.D
    bounds = [(0, D), (-0.25, 0.25)] * coord_count
    lim = np.array(bounds, dtype="float").T
    a1, a2 = 0.5 * (lim[0] + lim[1]), np.fabs(lim[0] - lim[1])
    G = lim.shape[1]
    n = max(5, max(1, popsize // len(bounds)) * G)
    s = attack.ReferenceStream(seed0 + p, n, G)
    left, right = pool.table[pool.ia[p]], pool.table[pool.ib[p]][None, :]

    def energy(genes):
        rows = attack.perturb(a1 + (genes - 0.5) * a2, left)
        out = oracle.score(rows, right, np.arange(len(rows), dtype=np.int32), np.zeros(len(rows), np.int32), pool.heads,
                           want_util=False, want_consensus=False)
        return 1 - out["per_learner"][:, 0, int(targets[p])]

    pop = s.init_population_lhs()
    pop, e = R.de_first(pop, energy(pop))
    traj = []
    for _ in range(maxiter):
        scale, r0, r1, cross = s.generation((0.5, 1.0), recombination)
        trial, flags = R.de_mutate(pop, scale, r0, r1, cross)
        trial = R.de_repair(trial, flags, s.repair(flags.sum()))
        pop, e = R.de_select(pop, e, trial, energy(trial))
        traj.append(e[0])
    return pop, e, np.array(traj, np.float32)


@pytest.mark.parametrize("scenario,problems", [("A", [0, 17, 63]), ("B", [0, 1, 2, 3, 4, 5, 6, 7])])
def test_host_stream_and_parallel_generation_equal_the_reference_solver(oracle, scenario, problems):
    fx = np.load(os.path.join(GOLD, "ref_de.npz"))
    pool, targets = mk.problem_inputs()
    np.testing.assert_array_equal(targets, fx["targets"])
    kw = {k: v for k, v in mk.SCENARIOS[scenario].items() if k != "problems"}
    for p in problems:
        pop, e, traj = run_numpy(oracle, pool, targets, p, **kw)
        j = mk.SCENARIOS[scenario]["problems"].index(p)
        np.testing.assert_array_equal(e.view(np.uint32), fx[f"{scenario}_energies"][j].view(np.uint32))
        np.testing.assert_array_equal(pop[0], fx[f"{scenario}_best"][j])
        assert hashlib.sha256(np.ascontiguousarray(pop).tobytes()).hexdigest() == str(fx[f"{scenario}_pop_sha"][j])
        np.testing.assert_array_equal(traj, fx[f"{scenario}_traj"][j])


def test_perturb_assigns_in_gene_order():
    emb = np.zeros(8, np.float32)
    out = attack.perturb(np.array([[1.9, 0.5, 1.2, -0.25, 7.99, 0.1, 9.0, 0.3]]), emb)
    np.testing.assert_array_equal(out[0], np.array([0, -0.25, 0, 0, 0, 0, 0, 0.3], np.float32))  # later gene wins; index clamps
