"""alink_b200.noise / Bagging.attackModel — the interface of /root/reference/code/noise.py:10-31,200-215 and
code/committee.py:28-37 on embedding pairs (host logic; the device-backed noise types are in tests/test_gpu_attack.py)."""
import numpy as np
import pytest

from alink_b200 import committee, noise


def _pairs(n=7, D=16, seed=0):
    rng = np.random.default_rng(seed)
    return [rng.normal(size=(n, D)).astype(np.float32), rng.normal(size=(n, D)).astype(np.float32)]


def test_attack_model_layout_is_lists_indexed_by_attack():
    X = _pairs()
    labels = np.eye(2, dtype=np.float32)[np.arange(7) % 2]
    # code/ALINK.py:93: every noise class is built as cls(model=..., sess=..., feature_model=...)
    attacks = [noise.get_relevant_noise(n)(model=None, sess=None, feature_model=None) for n in ("plain", "Gaussian", "speckle")]
    bag = committee.Bagging([], attacks)
    np.random.seed(3)
    out = bag.attackModel(X, 224, labels)  # target_size is an image size in the reference: ignored here
    assert len(out) == 2 and len(out[0]) == 3 and len(out[1]) == 3
    for side in (0, 1):
        np.testing.assert_array_equal(out[side][0], X[side])  # 'plain'
        for j in (1, 2):
            assert out[side][j].shape == X[side].shape and out[side][j].dtype == np.float32
            assert not np.array_equal(out[side][j], X[side])
    # the disguised-faces model is then fed [noisy[0][jj], noisy[1][jj]] per attack jj (ALINK.py:170)
    assert np.abs(out[0][1] - X[0]).max() < 0.1


def test_gaussian_consumes_the_stream_row_by_row_left_then_right():
    """code/noise.py:20-31: one np.random draw per row, all left rows first, then all right rows."""
    X = _pairs(5, 8)
    g = noise.Gaussian(mean=0.5, var=0.04)
    np.random.seed(11)
    left, right = g.addPairNoise(X, None)
    np.random.seed(11)
    dl = np.random.normal(0.5, 0.2, X[0].shape)  # row-major: identical to 5 draws of shape (8,)
    dr = np.random.normal(0.5, 0.2, X[1].shape)
    np.testing.assert_array_equal(left, (X[0] + dl).astype(np.float32))
    np.testing.assert_array_equal(right, (X[1] + dr).astype(np.float32))


def test_names_like_the_reference_mapping():
    assert noise.get_relevant_noise("PLAIN") is noise.Noise
    assert noise.get_relevant_noise("adversarial") is noise.AdversarialNoise
    assert noise.get_relevant_noise("coordinate") is noise.CoordinateNoise
    for name in ("saltpepper", "poisson", "perlin", "nonsense"):
        with pytest.raises(NotImplementedError):
            noise.get_relevant_noise(name)
