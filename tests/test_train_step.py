"""the finetune ("teach") minibatch step (code/siamese.py:52-58 per batch; ALINK.py:250).
CPU: the C oracle (fp32, fixed reduction orders) against an independent float64 numpy restatement of the Keras
math. GPU: csrc/train.cuh against the oracle, bit for bit, over several consecutive Adadelta steps."""
import numpy as np
import pytest

from alink_b200 import synth


def _setup(act, nb, seed=0, D=512, H1=512, H2=64):
    pool = synth.make_pool(1, N=max(nb, 64), M=400, K=1, act=act, D=D, H1=H1, H2=H2)
    rng = np.random.default_rng(seed)
    y = (rng.random(nb) < 0.5).astype(np.float32)
    params = [np.ascontiguousarray(a, np.float32).copy() for a in pool.heads.learner(0)]
    return pool, y, params


def _zeros_like(params):
    return [np.zeros_like(p) for p in params]


@pytest.mark.parametrize("act", ["softmax", "sigmoid"])
def test_oracle_train_step_matches_float64_keras_math(oracle, act):
    from oracle import ref_numpy
    nb = 16
    pool, y, params = _setup(act, nb)
    Cc = pool.heads.C
    ia, ib = pool.ia[:nb].copy(), pool.ib[:nb].copy()
    ag, ad = _zeros_like(params), _zeros_like(params)
    p64 = [p.astype(np.float64) for p in params]
    a64, d64 = [np.zeros_like(p) for p in p64], [np.zeros_like(p) for p in p64]
    for step in range(3):
        loss, acc, probs = oracle.train_step(pool.table, pool.table, ia, ib, y, params, ag, ad, 512, 512, 64, Cc,
                                             lr=0.1, want_probs=True)
        rl, ra, p64, a64, d64, rp = ref_numpy.train_step(pool.table[ia], pool.table[ib], y, p64, a64, d64, lr=0.1, act=act)
        assert abs(loss - rl) <= 1e-5 * max(1.0, abs(rl)), (step, loss, rl)
        assert acc == pytest.approx(ra, abs=1e-6)
        assert np.abs(probs - (rp if Cc == 2 else rp)).max() <= 2e-6
        for got, want in zip(params, p64):
            assert np.abs(got - want).max() <= 2e-5, step
    # the loss went down on the batch it was trained on
    l_after, _ = oracle.train_step(pool.table, pool.table, ia, ib, y, params, ag, ad, 512, 512, 64, Cc, apply=False)
    l_first, _ = oracle.train_step(pool.table, pool.table, ia, ib, y,
                                   [np.ascontiguousarray(a, np.float32).copy() for a in pool.heads.learner(0)],
                                   _zeros_like(params), _zeros_like(params), 512, 512, 64, Cc, apply=False)
    assert l_after < l_first


def test_oracle_evaluate_only_changes_nothing(oracle):
    pool, y, params = _setup("softmax", 8)
    before = [p.copy() for p in params]
    ag, ad = _zeros_like(params), _zeros_like(params)
    oracle.train_step(pool.table, pool.table, pool.ia[:8].copy(), pool.ib[:8].copy(), y, params, ag, ad, 512, 512, 64, 2,
                      apply=False)
    for a, b in zip(params, before):
        np.testing.assert_array_equal(a, b)
    assert all(not a.any() for a in ag + ad)


@pytest.mark.gpu
@pytest.mark.parametrize("act,nb,H1,H2", [("softmax", 16, 512, 64), ("sigmoid", 16, 512, 64), ("softmax", 1, 512, 64),
                                           ("softmax", 37, 128, 32)])
def test_gpu_train_steps_bit_identical_to_oracle(ctx, oracle, act, nb, H1, H2):
    from alink_b200 import _ffi
    pool, y, params = _setup(act, nb, H1=H1, H2=H2)
    Cc = pool.heads.C
    ctx.set_table(0, pool.table)
    ctx.set_heads(3, pool.heads, compute=_ffi.FP32_EXACT)
    ag, ad = _zeros_like(params), _zeros_like(params)
    shapes = [p.shape for p in params]
    for step in range(4):
        ia = pool.ia[step:step + nb].copy()
        ib = pool.ib[step:step + nb].copy()
        rl, ra, rp = oracle.train_step(pool.table, pool.table, ia, ib, y, params, ag, ad, 512, H1, H2, Cc, lr=0.1,
                                       want_probs=True)
        gl, ga, gp = ctx.train_step(3, 0, 0, 0, ia, ib, y, lr=0.1, want_probs=True, C_out=Cc)
        assert np.float32(gl).view(np.uint32) == np.float32(rl).view(np.uint32), (step, gl, rl)
        assert ga == ra
        np.testing.assert_array_equal(gp.view(np.uint32), rp.view(np.uint32))
        for which, ref in ((0, params), (1, ag), (2, ad)):
            got = ctx.get_train_state(3, 0, which, shapes)
            for a, b in zip(got, ref):
                np.testing.assert_array_equal(a.view(np.uint32), b.view(np.uint32))
    # scoring after training uses the updated device weights (exact head)
    per, _ = ctx.score_indexed(3, 0, 0, pool.ia[:64].copy(), pool.ib[:64].copy(), 64, 1, Cc)
    import copy
    heads = copy.deepcopy(pool.heads)
    heads.W1, heads.b1, heads.W2, heads.b2, heads.W3, heads.b3 = [[p] for p in params]
    ref = oracle.score(pool.table, pool.table, pool.ia[:64].copy(), pool.ib[:64].copy(), heads)
    np.testing.assert_array_equal(per.view(np.uint32), ref["per_learner"].view(np.uint32))


@pytest.mark.gpu
def test_finetune_like_alink_driver(ctx, oracle):
    """code/ALINK.py:250 `disguisedFacesModel.finetune([XL, XR], y, ft_epochs, 16, 1)`: every batch equals the
    oracle's step (same shuffles via np.random.seed), the loss goes down, the optimizer state persists, and the
    model then scores with the new weights."""
    from alink_b200 import siamese
    n = 200
    pool, _, params = _setup("softmax", n)
    ia, ib = pool.ia[:n], pool.ib[:n]
    XL, XR = pool.table[ia], pool.table[ib]
    y = np.asarray(pool.genuine[:n], dtype=np.float32)  # genuine / impostor labels of the synthetic pairs
    model = siamese.SiameseNetwork((512,), "unused", 0.1)
    model.set_weights(*params)
    np.random.seed(42)
    hist = model.finetune([XL, XR], y, 3, 16, 0)
    assert len(hist["loss"]) == 3 and hist["loss"][-1] < hist["loss"][0]
    # replay with the oracle: same split, same shuffles, same batches
    ag, ad = _zeros_like(params), _zeros_like(params)
    np.random.seed(42)
    split = int(n * 0.8)
    order = np.arange(split, dtype=np.int32)  # one index array shuffled in place every epoch (Keras _fit_loop)
    for epoch in range(3):
        np.random.shuffle(order)
        tot = 0.0
        for lo in range(0, split, 16):
            b = order[lo:lo + 16]
            l, _ = oracle.train_step(XL, XR, b, b, y[b], params, ag, ad, 512, 512, 64, 2, lr=0.1, eps=siamese.ADADELTA_EPS)
            tot += l * len(b)
        assert hist["loss"][epoch] == pytest.approx(tot / split, rel=1e-6)
    for got, want in zip(model.get_weights(), params):
        np.testing.assert_array_equal(got.view(np.uint32), want.view(np.uint32))
    for got, want in zip(model._opt[0], ag):
        np.testing.assert_array_equal(got, want)
    out = model.predict([XL[:32], XR[:32]])
    assert out.shape == (32, 2) and np.allclose(out.sum(1), 1, atol=1e-6)
    acc = model.testAccuracy(pool.table[:40], np.arange(40) // 4)
    assert 0.0 <= acc <= 1.0


@pytest.mark.gpu
def test_teach_runs_the_finetune_step_like_existing_al():
    """`learner.query(...)` then `learner.teach(X=..., y=..., only_new=True, verbose=1, epochs=2,
    validation_split=0.1)` exactly as /root/reference/code/existing_al.py:95-117 calls them: teach ==
    KerasClassifier.fit == Keras `Model.fit` (batch 32, no callbacks) on the head, bit-identical to driving
    `finetune`'s step loop by hand with the same shuffles; the known-data bookkeeping follows code/base.py:62-84."""
    from alink_b200 import engine, siamese, synth, uncertainty
    from alink_b200.keras_wrapper import KerasClassifier
    from alink_b200.learners import ActiveLearner, Committee
    engine.set_compute_mode("fp32")
    pool = synth.make_pool(1, N=400)
    XL, XR = pool.table[pool.ia], pool.table[pool.ib]
    Y = pool.genuine.astype(np.int64).reshape(-1, 1)

    def build(seed_name):
        net = siamese.SiameseNetwork((pool.D,), seed_name, 0.1)
        net.set_weights(*pool.heads.learner(0))
        return net

    net, twin = build("a"), build("b")
    learner = ActiveLearner(estimator=KerasClassifier(lambda: net), query_strategy=uncertainty.uncertainty_sampling, verbose=1)
    idx, inst = learner.query([XL, XR], n_instances=100, verbose=0)
    assert len(idx) == 100 and learner.X_training is None
    np.random.seed(11)
    learner.teach(X=[XL[idx], XR[idx]], y=Y[idx], only_new=True, verbose=0, epochs=2, validation_split=0.1)
    assert learner.estimator.classes_.tolist() == [0, 1]
    assert len(learner.X_training[0]) == 100 and len(learner.y_training) == 100
    np.random.seed(11)
    twin.finetune([XL[idx], XR[idx]], Y[idx], 2, 32, verbose=0, validation_split=0.1, _callbacks=False)
    for a, b in zip(net.get_weights(), twin.get_weights()):
        np.testing.assert_array_equal(a.view(np.uint32), b.view(np.uint32))
    assert any(not np.array_equal(a, b) for a, b in zip(net.get_weights(), pool.heads.learner(0)))  # it did train
    # a second batch of another size: each side is stacked with itself (base.py docstring)
    learner.teach(X=[XL[:37], XR[:37]], y=Y[:37], only_new=True, verbose=0)
    assert learner.X_training[0].shape == (137, pool.D) and learner.X_training[1].shape == (137, pool.D)
    assert 0.0 <= learner.score([XL, XR], Y) <= 1.0
    with pytest.raises(ValueError, match="inconsistent numbers of samples"):
        learner.teach(X=[XL[:5], XR[:5]], y=Y[:4])
    # committee: every learner is taught; rebag refits each on its own bootstrap of the known data
    nets = [build("c"), build("d")]
    com = Committee([ActiveLearner(KerasClassifier(lambda n=n: n)) for n in nets])
    com.teach([XL[:64], XR[:64]], Y[:64], verbose=0)
    assert com.classes_.tolist() == [0, 1]
    np.random.seed(3)
    com.rebag(verbose=0)
    assert not np.array_equal(nets[0].get_weights()[0], nets[1].get_weights()[0])  # different bootstraps
    p = com.predict_proba([XL, XR])
    assert p.shape == (400, 2)


@pytest.mark.gpu
@pytest.mark.parametrize("act", ["softmax", "sigmoid"])
def test_train_epoch_equals_the_step_loop_bit_for_bit(ctx, act):
    """alink_train_epoch = the batches' alink_train_steps enqueued back to back (one host round trip per pass instead of
    one per batch): per-batch losses / accuracies, parameters and both Adadelta accumulators must be bit-identical,
    with a ragged last batch, for the training pass and the evaluation pass (apply = 0)."""
    import time
    from alink_b200 import _ffi
    n, batch = 203, 16
    pool, _, params = _setup(act, n)
    y = np.asarray(pool.genuine[:n], dtype=np.float32)
    ia, ib = pool.ia[:n].copy(), pool.ib[:n].copy()
    shapes = [p.shape for p in params]
    ctx.set_table(0, pool.table)
    state = {}
    for how in ("steps", "epoch"):
        ctx.set_heads(3, pool.heads, compute=_ffi.FP32_EXACT)  # fresh weights, zero accumulators
        t0 = time.perf_counter()
        for _ in range(2):
            if how == "steps":
                la = [ctx.train_step(3, 0, 0, 0, ia[lo:lo + batch].copy(), ib[lo:lo + batch].copy(), y[lo:lo + batch], lr=0.1)
                      for lo in range(0, n, batch)]
                ls, accs = np.float32([v[0] for v in la]), np.float32([v[1] for v in la])
            else:
                ls, accs = ctx.train_epoch(3, 0, 0, 0, ia, ib, y, batch, lr=0.1)
        dt = (time.perf_counter() - t0) / 2
        if how == "steps":
            ev = [ctx.train_step(3, 0, 0, 0, ia[lo:lo + batch].copy(), ib[lo:lo + batch].copy(), y[lo:lo + batch], apply=False)
                  for lo in range(0, n, batch)]
            evl = np.float32([v[0] for v in ev])
        else:
            evl, _ = ctx.train_epoch(3, 0, 0, 0, ia, ib, y, batch, apply=False)
        state[how] = (ls, accs, evl, [ctx.get_train_state(3, 0, w, shapes) for w in (0, 1, 2)])
        print(f"{how}: {dt * 1e3:.2f} ms per pass of {len(ls)} batches ({dt * 1e6 / len(ls):.0f} us per batch)")
    a, b = state["steps"], state["epoch"]
    assert len(b[0]) == (n + batch - 1) // batch
    for i in range(3):
        np.testing.assert_array_equal(a[i].view(np.uint32), b[i].view(np.uint32))
    for wa, wb in zip(a[3], b[3]):
        for x, z in zip(wa, wb):
            np.testing.assert_array_equal(x.view(np.uint32), z.view(np.uint32))
