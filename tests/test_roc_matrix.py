"""the all-pairs score matrix (utilities/generateMatrixDFW.py) and the verification-curve
counts on it (utilities/ROC_precompute.py). CPU: the C oracle against the numpy restatement of the reference
loops. GPU: csrc/roc.cuh and alink_b200.matrix against the oracle, bit for bit (integer counts / fp32 scores)."""
import numpy as np
import pytest

from alink_b200 import synth


def _case(n, nthr, seed=5):
    rng = np.random.default_rng(seed)
    scores = rng.random((n, n)).astype(np.float32)
    mask = rng.integers(0, 6, size=(n, n)).astype(np.int32)  # codes 0 and 5 are ignored by every protocol
    thr = np.sort(rng.random(nthr))[::-1].copy()              # descending, like a thresholds.txt sweep
    thr[: min(5, nthr)] = scores[0, 1: 1 + min(5, nthr)].astype(np.float64)  # thresholds equal to scores: >= matters
    scores[3, 4] = np.nan
    return scores, mask, thr


@pytest.mark.parametrize("roc_case", [1, 2, 3])
def test_oracle_roc_counts_match_reference_loops(oracle, roc_case):
    from oracle import ref_numpy
    scores, mask, thr = _case(83, 41)
    tp, fp, ng, ni = oracle.roc_counts(scores, mask, thr, roc_case)
    tpr, fpr = ref_numpy.roc_precompute(scores, mask, thr, roc_case)
    np.testing.assert_array_equal(tp.astype(np.float64) / ng, tpr)
    np.testing.assert_array_equal(fp.astype(np.float64) / ni, fpr)


def test_oracle_roc_blocks_add_up(oracle):
    scores, mask, thr = _case(64, 17)
    whole = oracle.roc_counts(scores, mask, thr, 3)
    a = oracle.roc_counts(scores[:20], mask[:20], thr, 3, row0=0)
    b = oracle.roc_counts(scores[20:], mask[20:], thr, 3, row0=20)
    np.testing.assert_array_equal(a[0] + b[0], whole[0])
    np.testing.assert_array_equal(a[1] + b[1], whole[1])
    assert a[2] + b[2] == whole[2] and a[3] + b[3] == whole[3]


@pytest.mark.gpu
@pytest.mark.parametrize("roc_case", [1, 2, 3])
@pytest.mark.parametrize("n,nthr", [(1, 3), (257, 0), (700, 1000), (1500, 2047)])
def test_gpu_roc_counts_bit_exact(ctx, oracle, roc_case, n, nthr):
    scores, mask, thr = _case(n, nthr) if n > 8 else (np.zeros((n, n), np.float32), np.ones((n, n), np.int32), np.array([0.0, 1.0, -1.0])[:nthr])
    tp, fp, ng, ni = ctx.roc_counts(scores, mask, thr, roc_case)
    rtp, rfp, rng_, rni = oracle.roc_counts(scores, mask, thr, roc_case)
    np.testing.assert_array_equal(tp, rtp)
    np.testing.assert_array_equal(fp, rfp)
    assert (ng, ni) == (rng_, rni)


@pytest.mark.gpu
def test_gpu_roc_device_resident_blocks_add_up(ctx, oracle):
    import torch
    scores, mask, thr = _case(900, 300)
    whole = oracle.roc_counts(scores, mask, thr, 3)
    ds, dm = torch.from_numpy(scores).cuda(), torch.from_numpy(mask).cuda()
    tp = np.zeros(len(thr), np.uint64); fp = np.zeros(len(thr), np.uint64); ng = ni = 0
    for r0 in range(0, 900, 256):  # probe-row blocks, as a sharded all-pairs run would hold them
        t, f, g, i = ctx.roc_counts(ds[r0:r0 + 256], dm[r0:r0 + 256], thr, 3, row0=r0)
        tp += t; fp += f; ng += g; ni += i
    np.testing.assert_array_equal(tp, whole[0])
    np.testing.assert_array_equal(fp, whole[1])
    assert (ng, ni) == (whole[2], whole[3])


@pytest.mark.gpu
def test_gpu_roc_precompute_and_score_matrix_like_the_reference_scripts(ctx, oracle, tmp_path):
    """generateMatrixDFW.py -> ROC_precompute.py end to end on a 300-row table: matrix rows equal the per-row
    predict calls of the reference loop; the curve equals the numpy restatement on that matrix."""
    from alink_b200 import matrix, roc, siamese
    from oracle import ref_numpy
    pool = synth.make_pool(1, N=64, M=300, K=1)
    W1, b1, W2, b2, W3, b3 = pool.heads.learner(0)
    model = siamese.SiameseNetwork((512,), str(tmp_path / "m"), 0.1)
    model.set_weights(W1, b1, W2, b2, W3, b3)
    feats = pool.table
    np.save(tmp_path / "processedData.npy", feats)
    scores = matrix.generate_matrix(model, str(tmp_path / "processedData.npy"), str(tmp_path / "scores.txt"))
    assert scores.shape == (300, 300)
    for i in (0, 7, 299):  # generateMatrixDFW.py:28-33: X_left = features[i] repeated, X_right = all rows
        row = model.predict([np.repeat(feats[i:i + 1], 300, axis=0), feats])[:, 0]
        np.testing.assert_array_equal(scores[i].view(np.uint32), row.view(np.uint32))
    back = np.loadtxt(tmp_path / "scores.txt", dtype=float)  # ROC_precompute.py:11
    np.testing.assert_array_equal(back.astype(np.float32), scores)
    rng = np.random.default_rng(1)
    mask = rng.integers(1, 5, size=(300, 300)).astype(np.int32)
    thr = np.linspace(0, 1, 101)
    curve = roc.roc_precompute(scores, mask, thr, 3)
    tpr, fpr = ref_numpy.roc_precompute(back, mask, thr, 3)
    np.testing.assert_array_equal(curve[0], tpr)
    np.testing.assert_array_equal(curve[1], fpr)
