"""Pair enumeration order: index arrays vs the reference's nested loops
(/root/reference/code/readDFW.py:222-244, code/readMTP.py:123-135, restated in oracle/ref_numpy.py)."""
import numpy as np
import pytest

from alink_b200 import pairs
from oracle import ref_numpy


def _groups(rng, sizes, D=6):
    return [rng.standard_normal((n, D)).astype(np.float32) for n in sizes]


@pytest.mark.parametrize("ps,ds", [([3, 1, 2], [2, 4, 1]), ([1], [1]), ([2, 0, 3], [0, 2, 2]), ([4, 4], [1, 1, 1, 5]),
                                   ([0, 0], [3])])
def test_dfw_order_matches_reference_loops(ps, ds):
    rng = np.random.default_rng(len(ps) * 100 + len(ds))
    P, G = _groups(rng, ps), _groups(rng, ds)
    (XL, XR), Y = ref_numpy.create_mini_batch_dfw(P, G)
    ia, ib, y = pairs.pair_indices_dfw(ps, ds)
    assert ia.dtype == np.int32 and ib.dtype == np.int32
    table = np.concatenate(P + G)
    L, R = pairs.materialise(table, ia, ib)
    np.testing.assert_array_equal(L, XL)
    np.testing.assert_array_equal(R, XR)
    np.testing.assert_array_equal(y, Y)
    assert y.shape == Y.shape == (len(ia), 1)


@pytest.mark.parametrize("ps", [[3, 1, 2], [1], [2, 0, 3], [5, 5, 5, 5]])
def test_mtp_order_matches_reference_loops(ps):
    rng = np.random.default_rng(sum(ps))
    P = _groups(rng, ps)
    (XL, XR), Y = ref_numpy.create_mini_batch_mtp(P)
    ia, ib, y = pairs.pair_indices_mtp(ps)
    L, R = pairs.materialise(np.concatenate(P), ia, ib)
    np.testing.assert_array_equal(L, XL)
    np.testing.assert_array_equal(R, XR)
    np.testing.assert_array_equal(y, Y)


def test_empty_pool_raises_like_np_stack():
    with pytest.raises(ValueError, match="at least one array"):
        pairs.pair_indices_dfw([0, 0], [0])
    with pytest.raises(ValueError, match="at least one array"):
        ref_numpy.create_mini_batch_dfw([np.zeros((0, 4))], [np.zeros((0, 4))])


def test_large_pool_is_vectorised():
    ia, ib, y = pairs.pair_indices_dfw([20] * 60, [30] * 60)  # 60*60*600 + 60*60*900 = 5.4 M pairs
    assert len(ia) == 60 * 60 * (600 + 900)
    assert int(y.sum()) == 60 * 600 + 60 * 900
    assert ia.max() < 60 * 50 and ib.min() >= 60 * 20


@pytest.mark.gpu
def test_pool_scores_equal_materialised_scores():
    """`bag.predict(createMiniBatch(...))` (code/ALINK.py:151-160) through the index-driven pool equals the
    same call on the reference's materialised arrays, bit for bit."""
    from alink_b200 import committee, engine, siamese, synth
    engine.set_compute_mode("fp32")
    rng = np.random.default_rng(5)
    P, G = _groups(rng, [3, 2, 4], 512), _groups(rng, [2, 5, 1], 512)
    heads = synth.make_pool(1, N=64).heads
    nets = []
    for k in range(heads.K):
        n = siamese.SiameseNetwork((512,), f"m{k}", 0.1)
        n.set_weights(*heads.learner(k))
        nets.append(n)
    bag = committee.Bagging(nets, [])
    pool, Y = pairs.createMiniBatch(P, G)
    X, Yref = ref_numpy.create_mini_batch_dfw(P, G)
    np.testing.assert_array_equal(Y, Yref)
    np.testing.assert_array_equal(bag.predict(pool), bag.predict(X))
    pool2, Y2 = pairs.createMiniBatch(P)
    X2, _ = ref_numpy.create_mini_batch_mtp(P)
    np.testing.assert_array_equal(nets[0].predict(pool2), nets[0].predict(X2))


def test_property_random_ragged_groups():
    """hypothesis: any ragged group sizes (zeros included) give the loops' order."""
    from hypothesis import given, settings, strategies as st

    sizes = st.lists(st.integers(0, 4), min_size=1, max_size=5)

    @settings(max_examples=60, deadline=None)
    @given(ps=sizes, ds=sizes)
    def check(ps, ds):
        rng = np.random.default_rng(1)
        P, G = _groups(rng, ps, 3), _groups(rng, ds, 3)
        total = sum(ps) * sum(ds) + sum(ds) ** 2
        if total == 0:
            with pytest.raises(ValueError):
                pairs.pair_indices_dfw(ps, ds)
            return
        (XL, XR), Y = ref_numpy.create_mini_batch_dfw(P, G)
        ia, ib, y = pairs.pair_indices_dfw(ps, ds)
        L, R = pairs.materialise(np.concatenate(P + G), ia, ib)
        assert len(ia) == total
        np.testing.assert_array_equal(L, XL)
        np.testing.assert_array_equal(R, XR)
        np.testing.assert_array_equal(y, Y)

    check()


def test_split_disguise_data_like_reference():
    """code/readDFW.py:212-219."""
    from alink_b200 import readDFW
    rng = np.random.default_rng(0)
    G = _groups(rng, [5, 1, 4, 0])
    pre, post = readDFW.splitDisguiseData(G, pre_ratio=0.5)
    assert [len(g) for g in pre] == [2, 0, 2, 0] and [len(g) for g in post] == [3, 1, 2, 0]
    for g, a, b in zip(G, pre, post):
        np.testing.assert_array_equal(np.concatenate([a, b]), g)
    import inspect
    from alink_b200 import readMTP
    assert list(inspect.signature(readDFW.createMiniBatch).parameters)[:2] == ["X_plain", "X_dig"]
    assert list(inspect.signature(readMTP.createMiniBatch).parameters)[:1] == ["X_plain"]
