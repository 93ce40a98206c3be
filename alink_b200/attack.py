"""Embedding-space analogue of /root/reference/code/attack.py — the adversarial pool generator.

The reference's `PixelAttacker.attack_all(input_data, targets, dimensions, pixel_count=40, maxiter=50, popsize=250)`
(:91-103) runs, per image, a differential evolution over `pixel_count` (x, y, r, g, b) genes whose objective scores the
WHOLE population in one batched model call (code/differential_evolution.py:630-645, :692-714; `best1bin`, dither
[0.5, 1), `recombination=1, atol=-1, polish=False`, attack.py:79-81) and stops when the callback sees the target class
win (:64-65, :45-57). The image model (conv encoder) is upstream of this package's pool of embeddings, so the analogue
here perturbs `coord_count` COORDINATES of the left embedding of a pair: genes come in (coordinate index, value) pairs,
`perturb` assigns `e[int(index)] = value` the way `perturb_image` (:5-29) assigns `img[x, y] = rgb`, and the objective
is `1 - P(target class)` of the siamese head (:36-41 with `minimize=False`).

`EmbeddingAttacker.attack_all` keeps the reference's call shape. All pairs evolve in lockstep on the GPU (csrc/de.cuh):
one DE problem per pair, population arithmetic in float64, energies from the head kernel. Two random streams:
  * rng="reference": numpy RandomState per pair, drawn on the host in EXACTLY the reference solver's order (latin
    hypercube :433-469, dither :690-691, per member randint / shuffle / rand :782-808, constraint repair :775-780), so
    populations and energies equal the reference solver's generation by generation (tests/test_gpu_attack.py against a
    fixture produced by the reference's own differential_evolution.py);
  * rng="device": a counter-based stream on the GPU, no host round trip per generation (the production mode).
"""
from __future__ import annotations

import numpy as np

from . import _ffi, engine


def perturb(xs, emb):
    """perturb_image (code/attack.py:5-29) in embedding space: xs (n, 2 * coord_count) of (index, value) genes ->
    (n, D) copies of `emb` with emb[int(index)] = value, later genes overwriting earlier ones."""
    xs = np.asarray(xs)
    if xs.ndim < 2:
        xs = np.array([xs])
    out = np.tile(np.asarray(emb, dtype=np.float32), (len(xs), 1))
    D = out.shape[1]
    for x, row in zip(xs, out):
        for idx, val in np.split(x, len(x) // 2):
            row[min(max(int(idx), 0), D - 1)] = np.float32(val)
    return out


class ReferenceStream:
    """The reference solver's random stream for ONE problem (numpy RandomState, same calls in the same order)."""

    def __init__(self, seed, n, G):
        self.rng = np.random.RandomState(seed)
        self.n, self.G = n, G

    def init_population_lhs(self):
        rng, n, G = self.rng, self.n, self.G  # code/differential_evolution.py:433-469
        segsize = 1.0 / n
        samples = segsize * rng.random_sample((n, G)) + np.linspace(0., 1., n, endpoint=False)[:, np.newaxis]
        pop = np.zeros_like(samples)
        for j in range(G):
            order = rng.permutation(range(n))
            pop[:, j] = samples[order, j]
        return pop

    def generation(self, dither, recombination):
        """(scale, r0[n], r1[n], cross[n][G] or None) of one generation (:688-695, :782-808, :879-889)."""
        rng, n, G = self.rng, self.n, self.G
        scale = rng.rand() * (dither[1] - dither[0]) + dither[0]
        r0, r1 = np.empty(n, np.int32), np.empty(n, np.int32)
        cross = None if recombination >= 1.0 else np.empty((n, G), np.uint8)
        for c in range(n):
            fill = rng.randint(0, G)
            idxs = list(range(n))
            idxs.remove(c)
            rng.shuffle(idxs)
            r0[c], r1[c] = idxs[0], idxs[1]
            cr = rng.rand(G) < recombination
            cr[fill] = True
            if cross is not None:
                cross[c] = cr
        return scale, r0, r1, cross

    def repair(self, count):
        return np.array([self.rng.rand() for _ in range(int(count))], dtype=np.float64)  # :775-780, one draw per gene


class BatchedDE:
    """P differential evolutions in lockstep (one per pair) — the solver of code/differential_evolution.py:217-760 with
    its generation loop on the GPU. `specs`: the head(s) whose mean probability is the objective."""

    def __init__(self, specs, left, right, il, ir, targets, bounds, popsize=15, dither=(0.5, 1.0), recombination=0.7,
                 seed=None, rng="device", ctx=None):
        self.ctx = ctx or _ffi.default_context()
        self.specs = list(specs)
        self.K, self.C = len(self.specs), self.specs[0].C
        left = np.ascontiguousarray(left, dtype=np.float32)
        right = left if right is None else np.ascontiguousarray(right, dtype=np.float32)
        self.D = left.shape[1]
        self.slot_l, self.slot_r = _ffi.MAX_SLOTS - 2, _ffi.MAX_SLOTS - 1  # the dense staging slots
        self.ctx.set_table(self.slot_l, left)
        self.ctx.set_table(self.slot_r, right)
        self.il, self.ir = np.asarray(il, np.int32), np.asarray(ir, np.int32)
        self.targets = np.asarray(targets, np.int32)
        self.P = len(self.il)
        lim = np.array(bounds, dtype="float").T  # (2, G), :384
        self.lo, self.hi = np.ascontiguousarray(lim[0]), np.ascontiguousarray(lim[1])
        self.G = lim.shape[1]
        self.n = max(5, popsize * self.G)       # :415
        self.dither, self.cr = sorted(dither), recombination
        self.mode = rng
        base = 0 if seed is None else int(seed)
        self.seed = base
        if rng == "reference":
            self.streams = [ReferenceStream(base + p, self.n, self.G) for p in range(self.P)]
            pop = np.stack([s.init_population_lhs() for s in self.streams])
        else:
            g = np.random.Generator(np.random.PCG64(base))  # latin hypercube per problem, vectorised
            seg = 1.0 / self.n
            samples = seg * g.random((self.P, self.n, self.G)) + np.linspace(0., 1., self.n, endpoint=False)[None, :, None]
            order = np.argsort(g.random((self.P, self.n, self.G)), axis=1)
            pop = np.take_along_axis(samples, order, axis=1)
        self.ctx.de_init(pop, self.slot_l, self.slot_r, self.il, self.ir, self.targets, self.lo, self.hi)
        self.group = engine._GroupCache.group_for(self.ctx, self.specs)
        self.ctx.de_evaluate(self.group, 0)     # _calculate_population_energies, :630-668
        self.active = np.ones(self.P, np.uint8)
        self.nit = 0

    def step(self):
        """One generation for every active problem (`__next__`, :673-748)."""
        ctx = self.ctx
        self.nit += 1
        if self.mode == "reference":
            sc, r0, r1 = np.zeros(self.P), np.zeros((self.P, self.n), np.int32), np.ones((self.P, self.n), np.int32)
            cross = None if self.cr >= 1.0 else np.ones((self.P, self.n, self.G), np.uint8)
            for p, s in enumerate(self.streams):
                if self.active[p]:  # a stopped solver draws nothing more
                    sc[p], r0[p], r1[p], cr = s.generation(self.dither, self.cr)
                    if cross is not None:
                        cross[p] = cr
            nv = ctx.de_mutate_reference(sc, r0, r1, cross)  # a frozen problem (scale 0) yields pop[0]: never out of range
            off = np.concatenate([[0], np.cumsum(nv.astype(np.int64))])
            parts = [s.repair(nv[p]) for p, s in enumerate(self.streams) if nv[p]]
            ctx.de_repair(np.concatenate(parts) if parts else np.zeros(0), off)
        else:
            ctx.de_mutate_device(self.seed, self.nit, self.dither, self.cr)
        ctx.de_evaluate(self.group, 1)
        ctx.de_select(self.active if not self.active.all() else None)

    def best(self, want_rows=True):
        """(per-learner probabilities (P, K, C) of the best members, their perturbed left rows (P, D))."""
        return self.ctx.de_evaluate(self.group, 2, K=self.K, Cc=self.C, D=self.D, want_rows=want_rows)

    def state(self):
        return self.ctx.de_get()


class EmbeddingAttacker:
    """`PixelAttacker` (code/attack.py:32-103) for embedding pairs. `model`: a SiameseNetwork, a Bagging / Committee, or
    anything exposing `_alink_specs()`."""

    def __init__(self, model, ctx=None):
        self.model = model
        self.ctx = ctx
        self.specs = engine.specs_of(model)
        if self.specs is None:
            raise TypeError("EmbeddingAttacker needs alink_b200 models (no host fallback)")

    def attack_all(self, input_data, targets, dimensions=None, coord_count=40, maxiter=50, popsize=250, verbose=False,
                   value_range=(-0.25, 0.25), seed=None, rng="device", check_every=1, return_solver=False):
        """input_data = [X_left (N, D), X_right (N, D)]; targets (N, 2) one-hot (or (N,) class ids) of the class each
        attack should REACH (code/attack.py:96-98: `target_class = np.argmax(targets[i])`, a targeted attack).
        Returns the perturbed X_left (N, D). `dimensions` (the image size of the reference) defaults to (D,)."""
        XL, XR = engine._as_pair(input_data)
        N, D = XL.shape
        t = np.asarray(targets)
        tc = np.argmax(t, axis=1).astype(np.int32) if t.ndim == 2 else t.astype(np.int32)
        dim = D if dimensions is None else int(np.prod(dimensions))
        bounds = [(0, dim), (float(value_range[0]), float(value_range[1]))] * coord_count      # :71
        popmul = max(1, popsize // len(bounds))                                                  # :74
        de = BatchedDE(self.specs, XL, XR, np.arange(N), np.arange(N), tc, bounds, popsize=popmul, dither=(0.5, 1.0),
                       recombination=1, seed=seed, rng=rng, ctx=self.ctx)
        for it in range(1, maxiter + 1):
            de.step()
            if check_every and it % check_every == 0:
                probs, _ = de.best(want_rows=False)
                mean = probs.astype(np.float64).mean(axis=1) if de.K > 1 else probs[:, 0, :]
                if de.C == 1:
                    pred = (mean[:, 0] > 0.5).astype(np.int32)
                else:
                    pred = np.argmax(mean, axis=1).astype(np.int32)
                de.active &= (pred != tc).astype(np.uint8)  # attack_success (:45-57): stop once the target class wins
                if verbose:
                    print("generation %d: %d of %d attacks succeeded" % (it, int((de.active == 0).sum()), N))
                if not de.active.any():
                    break
        _, rows = de.best(want_rows=True)
        return (rows, de) if return_solver else rows

    def attack(self, pair, actual_class, target, coord_count, dimensions=None, maxiter=75, popsize=400, verbose=False, **kw):
        """One pair (code/attack.py:59-89). `target` None = untargeted: reach the class that is not `actual_class`."""
        l, r = np.asarray(pair[0], np.float32).reshape(1, -1), np.asarray(pair[1], np.float32).reshape(1, -1)
        t = np.array([1 - actual_class if target is None else target], np.int32)
        return self.attack_all([l, r], t, dimensions, coord_count, maxiter, popsize, verbose, **kw)[0]
