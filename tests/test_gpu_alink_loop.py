"""The A-LINK loop of /root/reference/code/ALINK.py:146-255 on embedding pairs, written with the drop-in modules the way the
driver writes it — `noise.get_relevant_noise(x)(model=M2, ...)`, `committee.Bagging(ensemble, ensembleNoise)`, `bag.predict`,
`bag.attackModel`, `M2.predict` per noise, the rule, the per-noise partition, `M2.finetune(..., ft_epochs, 16, 1)` — and every
iteration checked against the CPU oracle: committee scores bit for bit, the rule's selection and ACTIVE_COUNT, the pairs
handed to the finetune, and that the finetuned model is the one that scores the next batch."""
import numpy as np
import pytest

from alink_b200 import alink, committee, engine, noise, siamese, synth

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _fp32_mode():
    engine.set_compute_mode("fp32")
    yield
    engine.set_compute_mode("fp32")


def test_three_iterations_of_the_loop_against_the_oracle(oracle):
    from oracle import ref_numpy as R
    pool = synth.make_pool(1, N=1800, M=900)
    nets = []
    for k in range(3):
        net = siamese.SiameseNetwork((pool.D,), "/tmp/alink_loop_model_%d" % k, 0.1)
        net.set_weights(*pool.heads.learner(k))
        nets.append(net)
    ensemble, m2 = nets[:2], nets[2]
    names = ["plain", "gaussian", "coordinate"]                                              # FLAGS.noise.split(',')
    ensembleNoise = [noise.get_relevant_noise(x)(model=m2, sess=None, feature_model=None) for x in names]   # ALINK.py:93
    ensembleNoise[1].var, ensembleNoise[1].sigma = 4e-4, 0.02
    bag = committee.Bagging(ensemble, ensembleNoise)                                         # :97
    state = alink.ALinkState()
    np.random.seed(7)
    bs = 600
    w_before = [w.copy() for w in m2.get_weights()]
    import copy
    heads12 = copy.deepcopy(pool.heads)
    for f in ("W1", "b1", "W2", "b2", "W3", "b3"):
        setattr(heads12, f, getattr(pool.heads, f)[:2])
    heads12.K = 2
    for it in range(3):
        lo = it * bs
        X = [np.ascontiguousarray(pool.table[pool.ia[lo:lo + bs]]), np.ascontiguousarray(pool.table[pool.ib[lo:lo + bs]])]
        y = np.asarray(pool.genuine[lo:lo + bs], dtype=np.float32).reshape(-1, 1)           # batch_y: the pseudo-oracle
        # replay of the random draws of this iteration's host noise ('gaussian': left rows, then right rows)
        rs = np.random.get_state()
        dl = np.random.normal(0.0, 0.02, X[0].shape)
        dr = np.random.normal(0.0, 0.02, X[1].shape)
        np.random.set_state(rs)
        m2_heads = copy.deepcopy(pool.heads)
        for f, w in zip(("W1", "b1", "W2", "b2", "W3", "b3"), m2.get_weights()):
            setattr(m2_heads, f, [w])
        m2_heads.K = 1
        pend0, ft0 = state.pending(), state.finetunes
        idx, ens = alink.alink_iteration(bag, m2, X, y, state, target_size=224, col=1, disparity_ratio=0.25, eps=0.05,
                                         batch_send=10, ft_epochs=2, verbose=0)
        # committee scores (:160)
        ref = oracle.score(X[0], X[1], None, None, heads12)
        np.testing.assert_array_equal(ens.view(np.uint32), ref["consensus"].view(np.uint32))
        # the three noisy copies, scored by the oracle's head with M2's CURRENT weights (:170)
        noisy = [(X[0], X[1]),
                 ((X[0] + dl).astype(np.float32), (X[1] + dr).astype(np.float32)),
                 (R.perturb_rows(X[0], np.arange(bs), 0, 0, 0), R.perturb_rows(X[1], np.arange(bs), 1, 0, 0))]
        p2 = np.stack([oracle.score(a, b, None, None, m2_heads)["per_learner"][:, 0, 1] for a, b in noisy])
        ridx, ractive = oracle.disparity_select(p2, ref["consensus"][:, 1], 0.25, 0.05, False, y[:, 0])
        np.testing.assert_array_equal(idx, ridx)
        assert len(idx) > 0
        mp = len(idx) // 3
        if state.finetunes == ft0:
            assert state.pending() == pend0 + 3 * mp                                         # :216-225
        else:
            assert state.pending() == 0 and pend0 + 3 * mp >= 10                             # :227, :247-250
    assert state.UN_SIZE == 3 * bs and state.ACTIVE_COUNT >= len(idx)
    assert state.finetunes >= 1
    assert any(not np.array_equal(a, b) for a, b in zip(w_before, m2.get_weights()))         # M2 was finetuned on the device
