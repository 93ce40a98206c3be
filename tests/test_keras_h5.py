"""Keras `.h5` weight files without h5py (/root/reference/code/siamese.py:114-125).

Pinning: `tests/golden/libhdf5_written_testdouble.mat` is a file written by a real libhdf5 (a MATLAB 7.4
`-v7.3` save; it is SciPy's own test fixture scipy/io/matlab/tests/data/testhdf5_7.4_GLNX86.mat, BSD-licensed,
whose expected content SciPy's test_mio.py states: `testdouble` = 0, pi/4, ..., 2 pi). It exercises the
superblock behind a 512-byte user block, the symbol-table group (B-tree v1, local heap, SNOD), a version-1
object header, dataspace, IEEE-float datatype, a fixed-length string attribute and a contiguous layout.
The Keras tree itself is only checked against files from our own writer (no Keras-written file exists here).
"""
import itertools
import os

import numpy as np
import pytest

from alink_b200 import keras_h5 as kh

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _head(rng, D=32, H1=16, H2=8, C=2):
    return [rng.standard_normal(s).astype(np.float32) for s in [(D, H1), (H1,), (H1, H2), (H2,), (H2, C), (C,)]]


def test_reads_a_file_written_by_libhdf5():
    f = kh.H5File(os.path.join(GOLD, "libhdf5_written_testdouble.mat"))
    assert list(f.root.children) == ["testdouble"]
    d = f.root["testdouble"]
    assert d.attrs["MATLAB_class"] == b"double"
    assert d.value.dtype == np.float64 and d.value.shape == (9, 1)
    np.testing.assert_allclose(d.value.ravel(), np.arange(9) * np.pi / 4, rtol=0, atol=1e-15)


def test_head_roundtrip_reference_layer_names(tmp_path):
    arrs = _head(np.random.default_rng(0), 512, 512, 64, 2)
    p = str(tmp_path / "m.h5")
    kh.save_head_weights(p, arrs)
    tree = kh.read_keras_weights(p)
    assert [l for l, _ in tree] == ["input_1", "input_2", "lambda_1", "dense_1", "dense_2", "dense_3", "activation_1"]  # code/siamese.py:24-33
    assert [n for n, _ in tree[3][1]] == ["dense_1/kernel:0", "dense_1/bias:0"]
    f = kh.H5File(p)
    assert f.root.attrs["backend"] == b"tensorflow" and f.root.attrs["keras_version"] == b"2.1.2"
    assert f.root["dense_1/dense_1/kernel:0"].value.shape == (512, 512)  # the "/" in the name nests a group
    out = kh.load_head_weights(p, 512, 512, 64, 2)
    for a, b in zip(arrs, out):
        assert b.dtype == np.float32 and np.array_equal(a.view(np.uint32), b.view(np.uint32))


@pytest.mark.parametrize("sb,mw", list(itertools.product([0, 1, 2], [False, True])))
def test_structural_variants(tmp_path, sb, mw):
    arrs = _head(np.random.default_rng(1))
    layers = [("input_1", [])] + [(f"dense_{i}", [(f"dense_{i}/kernel:0", arrs[2 * i - 2]), (f"dense_{i}/bias:0", arrs[2 * i - 1])])
                                  for i in (1, 2, 3)]
    p = str(tmp_path / "v.h5")
    for av, sv, cont, comp, vl in itertools.product([1, 2, 3], [1, 2], [False, True], [False, True], [False, True]):
        kh.write_keras_weights(p, layers, model_weights_group=mw, attr_version=av, space_version=sv,
                               continuation=cont, compact=comp, vlen_names=vl, superblock=sb)
        got = kh.load_head_weights(p, 32, 16, 8, 2)
        for a, b in zip(arrs, got):
            np.testing.assert_array_equal(a, b)


def test_many_layers_span_several_symbol_nodes(tmp_path):
    rng = np.random.default_rng(2)
    layers = [(f"conv_{i}", [(f"conv_{i}/kernel:0", rng.standard_normal((3, 3, 2, 2)).astype(np.float32))]) for i in range(40)]
    arrs = _head(rng)
    layers += [(f"dense_{i}", [(f"dense_{i}/kernel:0", arrs[2 * i - 2]), (f"dense_{i}/bias:0", arrs[2 * i - 1])]) for i in (1, 2, 3)]
    p = str(tmp_path / "big.h5")
    kh.write_keras_weights(p, layers)
    tree = kh.read_keras_weights(p)
    assert [l for l, _ in tree] == [l for l, _ in layers]  # layer_names order, not the B-tree's sorted order
    for (_, w), (_, w2) in zip(layers, tree):
        np.testing.assert_array_equal(w[0][1], w2[0][1])
    got = kh.load_head_weights(p, 32, 16, 8, 2)  # the head is the last three (kernel, bias) pairs
    np.testing.assert_array_equal(got[0], arrs[0])


def test_errors(tmp_path):
    p = str(tmp_path / "x.h5")
    open(p, "wb").write(b"not hdf5" * 100)
    with pytest.raises(kh.H5FormatError, match="signature"):
        kh.H5File(p)
    arrs = _head(np.random.default_rng(3))
    data = kh.save_head_weights(p, arrs)
    with pytest.raises(kh.H5FormatError, match="do not match"):
        kh.load_head_weights(p, 64, 16, 8, 2)
    with pytest.raises(kh.H5FormatError, match="past the end"):
        kh.H5File(data[: len(data) // 2]).root  # truncated file
    kh.write_keras_weights(p, [("dense_1", [("dense_1/kernel:0", arrs[0])])])
    with pytest.raises(kh.H5FormatError, match="needs 6"):
        kh.load_head_weights(p, 32, 16, 8, 2)
    f = kh.H5File(os.path.join(GOLD, "libhdf5_written_testdouble.mat"))
    assert "layer_names" not in f.root.attrs
    with pytest.raises(kh.H5FormatError, match="layer_names"):
        kh.read_keras_weights(os.path.join(GOLD, "libhdf5_written_testdouble.mat"))


@pytest.mark.parametrize("mod", ["siamese", "siamese3"])
def test_shim_save_and_maybe_load(tmp_path, mod):
    """`save()` / `maybeLoadFromMemory()` of the drop-in classes (code/siamese.py:114-125, siamese3.py:88-99)."""
    import importlib
    m = importlib.import_module("alink_b200." + mod)
    name = str(tmp_path / "ensemble0")
    net = m.SiameseNetwork((512,), name, 0.1, seed=7)
    assert net.maybeLoadFromMemory() is False
    net.save()
    assert os.path.exists(name + ".h5")
    net2 = m.SiameseNetwork((512,), name, 0.1, seed=8)
    assert net2.maybeLoadFromMemory() is True
    for a, b in zip(net.get_weights(), net2.get_weights()):
        np.testing.assert_array_equal(a, b)
    wrong = m.SiameseNetwork((2048,), name, 0.1)  # shape mismatch: swallowed, like Keras' ValueError is
    assert wrong.maybeLoadFromMemory() is False
    os.remove(name + ".h5")  # the older .npz checkpoints still load
    W = net.get_weights()
    np.savez(name + ".npz", W1=W[0], b1=W[1], W2=W[2], b2=W[3], W3=W[4], b3=W[5])
    net3 = m.SiameseNetwork((512,), name, 0.1, seed=9)
    assert net3.maybeLoadFromMemory() is True
    np.testing.assert_array_equal(net3.get_weights()[4], W[4])
    net.save(customName=str(tmp_path / "other"))
    assert os.path.exists(str(tmp_path / "other.h5"))


def test_property_random_trees_and_dtypes(tmp_path):
    """hypothesis: random layer lists, weight shapes (rank 0-4, empty dims) and dtypes round-trip bit for bit
    in every writer variant."""
    from hypothesis import given, settings, strategies as st

    dtypes = st.sampled_from([np.float32, np.float64, np.float16, np.int32, np.int64, np.uint8])
    shape = st.lists(st.integers(0, 5), min_size=0, max_size=4).map(tuple)
    weight = st.tuples(shape, dtypes)
    layer = st.lists(weight, min_size=0, max_size=3)
    p = str(tmp_path / "h.h5")

    @settings(max_examples=40, deadline=None)
    @given(layers=st.lists(layer, min_size=1, max_size=12), av=st.sampled_from([1, 2, 3]), sv=st.sampled_from([1, 2]),
           cont=st.booleans(), comp=st.booleans(), vl=st.booleans(), sb=st.sampled_from([0, 1, 2]), seed=st.integers(0, 99),
           chunks=st.sampled_from([None, None, 2, 3]))
    def check(layers, av, sv, cont, comp, vl, sb, seed, chunks):
        rng = np.random.default_rng(seed)
        L = []
        for i, ws in enumerate(layers):
            name = f"layer_{i}"
            L.append((name, [(f"{name}/w{j}:0", (rng.standard_normal(s) * 50).astype(dt)) for j, (s, dt) in enumerate(ws)]))
        kh.write_keras_weights(p, L, attr_version=av, space_version=sv, continuation=cont, compact=comp,
                               vlen_names=vl, superblock=sb, chunks=chunks)
        got = kh.read_keras_weights(p)
        assert [n for n, _ in got] == [n for n, _ in L]
        for (_, w), (_, w2) in zip(L, got):
            assert [n for n, _ in w] == [n for n, _ in w2]
            for (_, a), (_, b) in zip(w, w2):
                assert a.dtype == b.dtype and a.shape == b.shape
                assert a.tobytes() == b.tobytes()

    check()


def test_chunked_unfiltered_datasets(tmp_path):
    arrs = _head(np.random.default_rng(4), 13, 7, 5, 2)
    layers = [(f"dense_{i}", [(f"dense_{i}/kernel:0", arrs[2 * i - 2]), (f"dense_{i}/bias:0", arrs[2 * i - 1])]) for i in (1, 2, 3)]
    p = str(tmp_path / "c.h5")
    kh.write_keras_weights(p, layers, chunks=4)  # 13 x 7 in 4 x 4 chunks: edge chunks are cropped
    for a, b in zip(arrs, kh.load_head_weights(p, 13, 7, 5, 2)):
        np.testing.assert_array_equal(a, b)
