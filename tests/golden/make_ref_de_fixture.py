"""Golden vectors for the adversarial pool generator, produced by RUNNING THE REFERENCE'S OWN SOLVER
(/root/reference/code/differential_evolution.py, the SciPy fork attack.py drives) in this container:
    python tests/golden/make_ref_de_fixture.py      -> tests/golden/ref_de.npz
The solver module is imported unmodified; it needs three names newer SciPy moved or dropped, which are shimmed:
`scipy.optimize.optimize._status_message` (a dict of four strings), `scipy._lib.six.xrange / string_types`.
Objective = `1 - P(target)` of the C oracle's siamese head (oracle/alink_oracle.c) on the embedding-space analogue of
perturb_image (attack.py:5-29): genes are (coordinate index, value) pairs, e[int(index)] = float32(value).
Scenario A = attack.py's settings (best1bin, dither [0.5, 1), recombination 1, atol -1, polish off, 40 coordinates,
popsize 250 -> 240 members) for 10 generations on 64 pairs; scenario B = recombination 0.7 (binomial crossover masks).
Stored per problem: final energies, best member, sha256 of the final population, best energy per generation."""
import hashlib
import importlib
import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = "/root/reference/code"
sys.path.insert(0, ROOT)


def import_reference_de():
    import scipy.optimize
    m = types.ModuleType("scipy.optimize.optimize")
    m._status_message = {"success": "Optimization terminated successfully.",
                         "maxfev": "Maximum number of function evaluations has been exceeded.",
                         "maxiter": "Maximum number of iterations has been exceeded.",
                         "pr_loss": "Desired error not necessarily achieved due to precision loss."}
    sys.modules["scipy.optimize.optimize"] = m
    six = types.ModuleType("scipy._lib.six")
    six.xrange, six.string_types = range, (str,)
    sys.modules["scipy._lib.six"] = six
    sys.path.insert(0, REF)
    try:
        return importlib.import_module("differential_evolution")
    finally:
        sys.path.remove(REF)


def problem_inputs(P=64):
    """64 pairs of configs[0]'s pool scored by ONE learner; the attack aims at the class the model does not predict."""
    from alink_b200 import synth
    from oracle import oracle_ffi
    oracle_ffi.build()
    pool = synth.make_pool(1, N=P, K=1)
    out = oracle_ffi.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads, want_util=False)
    targets = (1 - np.argmax(out["per_learner"][:, 0, :], axis=1)).astype(np.int32)
    return pool, targets


def perturb_rows(xs, emb):
    xs = np.asarray(xs)
    out = np.tile(emb, (len(xs), 1)).astype(np.float32)
    D = out.shape[1]
    for x, row in zip(xs, out):
        for j in range(0, len(x), 2):
            row[min(max(int(x[j]), 0), D - 1)] = np.float32(x[j + 1])
    return out


def run(de_mod, pool, targets, problems, coord_count, popsize, maxiter, recombination, seed0, value_range=(-0.25, 0.25)):
    from oracle import oracle_ffi
    D = pool.D
    bounds = [(0, D), value_range] * coord_count
    popmul = max(1, popsize // len(bounds))
    res = {"energies": [], "best": [], "pop_sha": [], "traj": []}
    for p in problems:
        left, right = pool.table[pool.ia[p]], pool.table[pool.ib[p]][None, :]
        tc = int(targets[p])

        def func(xs):
            rows = perturb_rows(xs, left)
            out = oracle_ffi.score(rows, right, np.arange(len(rows), dtype=np.int32), np.zeros(len(rows), np.int32),
                                   pool.heads, want_util=False, want_consensus=False)
            return 1 - out["per_learner"][:, 0, tc]

        traj = []
        solver = de_mod.DifferentialEvolutionSolver(func, bounds, maxiter=maxiter, popsize=popmul, recombination=recombination,
                                                    atol=-1, polish=False, seed=seed0 + p,
                                                    callback=lambda x, convergence: traj.append(np.float32(solver.population_energies[0])))
        solver.solve()
        res["energies"].append(np.asarray(solver.population_energies, dtype=np.float32))
        res["best"].append(np.asarray(solver.population[0], dtype=np.float64))
        res["pop_sha"].append(hashlib.sha256(np.ascontiguousarray(solver.population, dtype=np.float64).tobytes()).hexdigest())
        res["traj"].append(np.asarray(traj, dtype=np.float32))
    return {k: np.array(v) for k, v in res.items()}


SCENARIOS = {
    "A": dict(problems=list(range(64)), coord_count=40, popsize=250, maxiter=10, recombination=1, seed0=1000),
    "B": dict(problems=list(range(8)), coord_count=6, popsize=60, maxiter=8, recombination=0.7, seed0=2000),
}

if __name__ == "__main__":
    de_mod = import_reference_de()
    pool, targets = problem_inputs()
    out = {"targets": targets}
    for name, kw in SCENARIOS.items():
        r = run(de_mod, pool, targets, **kw)
        for k, v in r.items():
            out[f"{name}_{k}"] = v
        print(name, r["energies"].shape, r["traj"][:2])
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "ref_de.npz"), **out)
