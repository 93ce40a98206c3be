"""BASELINE configs[3] / [4] shaped cases at test size: the A2-LINK noise-augmented pool (disparity rule over
T perturbed copies, code/ALINK.py:160-201) and DFW-style probe x gallery all-pairs scoring
(utilities/generateMatrixDFW.py:27-35, code/ALINK_MTP.py:279-288), against the CPU oracle."""
import numpy as np
import pytest

from alink_b200 import _ffi, synth

pytestmark = pytest.mark.gpu


def _allpairs_indices(P, G, row0=0):
    i = np.repeat(np.arange(row0, row0 + P, dtype=np.int32), G)
    j = np.tile(np.arange(G, dtype=np.int32), P)
    return i, j


@pytest.mark.parametrize("mode", ["fp32", "refine", "bf16"])
def test_allpairs_query_corner_vs_oracle(ctx, oracle, mode):
    """A probe x gallery corner, K = 3, uncertainty ranking; pair index = i * G + j."""
    pool = synth.make_pool(1, N=64, M=1200)
    P, G, B = 160, 700, 200
    probe = np.ascontiguousarray(pool.table[:P + 40])
    gallery = np.ascontiguousarray(pool.table[300:300 + G])
    ctx.set_table(3, probe)
    ctx.set_table(4, gallery)
    ctx.set_heads(3, pool.heads, compute=_ffi.COMPUTE_ID[mode])
    idx, util, keys = ctx.query_allpairs(3, 3, 4, 0, P, "uncertainty", B)
    ia, ib = _allpairs_indices(P, G)
    ref = oracle.score(probe, gallery, ia, ib, pool.heads, measure="uncertainty")
    ridx, rutil, rkeys = oracle.topk(ref["util"], B)
    if mode == "bf16":
        assert np.abs(util - ref["util"][idx]).max() <= 2e-2
        overlap = len(set(idx.tolist()) & set(ridx.tolist())) / B  # reported, not an acceptance criterion
        print(f"bf16 all-pairs top-{B} overlap with fp32: {overlap:.2f}")
        assert overlap > 0.2
    else:
        np.testing.assert_array_equal(keys, rkeys)
        np.testing.assert_array_equal(idx, ridx)
        np.testing.assert_array_equal(util.view(np.uint32), rutil.view(np.uint32))
    if mode == "refine":
        assert ctx.refine_stats()[0] >= B


def test_allpairs_probe_block_shards_merge_to_single_shot(ctx, oracle):
    """Sharding by probe-row block: per-block top-B keys merged == one-shot top-B."""
    pool = synth.make_pool(1, N=64, M=900)
    P, G, B = 96, 500, 150
    probe, gallery = np.ascontiguousarray(pool.table[:P]), np.ascontiguousarray(pool.table[400:400 + G])
    ctx.set_table(3, probe)
    ctx.set_table(4, gallery)
    ctx.set_heads(3, pool.heads, compute=_ffi.FP32_EXACT)
    _, _, whole = ctx.query_allpairs(3, 3, 4, 0, P, "entropy", B)
    parts = []
    for r0, nr in ((0, 30), (30, 40), (70, 26)):
        _, _, k = ctx.query_allpairs(3, 3, 4, r0, nr, "entropy", B)
        parts.append(k)
    merged = ctx.topk_keys(np.concatenate(parts), B)
    np.testing.assert_array_equal(merged, whole)
    ia, ib = _allpairs_indices(P, G)
    ref = oracle.score(probe, gallery, ia, ib, pool.heads, measure="entropy")
    np.testing.assert_array_equal(whole, oracle.topk(ref["util"], B)[2])


def test_score_matrix_block_like_generate_matrix(ctx, oracle):
    """scores[i][j] = model.predict([features[i]] * G, features)[j][0] (utilities/generateMatrixDFW.py:28-35)."""
    pool = synth.make_pool(1, N=64, K=1, M=400)
    feats = np.ascontiguousarray(pool.table[:257])
    ctx.set_table(3, feats)
    ctx.set_heads(3, pool.heads, compute=_ffi.FP32_EXACT)
    _, cons = ctx.score_allpairs(3, 3, 3, 5, 64, len(feats), 1, 2)
    mat = cons[:, 0].reshape(64, len(feats))
    ia, ib = _allpairs_indices(64, len(feats), row0=5)
    ref = oracle.score(feats, feats, ia, ib, pool.heads)
    np.testing.assert_array_equal(mat, ref["per_learner"][:, 0, 0].reshape(64, len(feats)))


def test_allpairs_16m_pairs_properties(ctx):
    """4096 x 4096 = 16.8 M pairs in the tensor-core mode (size-independent properties): the selection is
    sorted by (utility desc, index asc), every index is inside the block, and the top-B of a sub-block
    that contains the winners is the same set."""
    pool = synth.make_pool(1, N=64, M=4096)
    ctx.set_table(3, pool.table)
    ctx.set_heads(3, pool.heads, compute=_ffi.BF16_TC)
    B = 1000
    idx, util, keys = ctx.query_allpairs(3, 3, 3, 0, 4096, "uncertainty", B)
    assert np.all(keys[:-1] > keys[1:])
    assert idx.min() >= 0 and idx.max() < 4096 * 4096 and len(set(idx.tolist())) == B
    rows = np.unique(idx // 4096)
    r0, r1 = int(rows.min()), int(rows.max()) + 1
    idx2, util2, keys2 = ctx.query_allpairs(3, 3, 3, r0, r1 - r0, "uncertainty", B)
    np.testing.assert_array_equal(keys2, keys)


@pytest.mark.parametrize("blind", [False, True])
def test_a2link_noise_pool_disparity_at_scale(ctx, oracle, blind):
    """configs[3] shape at 1/5 scale: 200 k base pairs x T = 8 perturbed copies through the disparity rule."""
    rng = np.random.default_rng(4)
    N, T = 200_000, 8
    p1 = rng.random(N).astype(np.float32)
    p2 = np.clip(p1[None, :] + rng.normal(0, 0.15, size=(T, N)), 0, 1).astype(np.float32)
    p2[:, ::1000] = p1[::1000]  # exact ties at zero disparity
    y = (rng.random(N) < 0.5).astype(np.float32)
    idx, active = ctx.disparity_select(p2, p1, 0.25, 0.05, blind, y)
    ridx, ractive = oracle.disparity_select(p2, p1, 0.25, 0.05, blind, y)
    np.testing.assert_array_equal(idx, ridx)
    assert active == ractive
    assert np.all(np.diff(idx) > 0)


@pytest.mark.parametrize("blind", [False, True])
@pytest.mark.parametrize("shards", [2, 3])
def test_sharded_disparity_equals_single_shot(ctx, oracle, blind, shards):
    """the per-noise prefix is a GLOBAL order statistic. Shards of one pool (here: in one process, the
    histograms summed by hand where dist.py all-reduces them) must select exactly the single-shot set."""
    import torch
    from alink_b200.dist import global_threshold_key, shard_range
    rng = np.random.default_rng(11)
    N, T = 150_001, 4
    p1 = rng.random(N).astype(np.float32)
    p2 = np.clip(p1[None, :] + 0.2 * rng.standard_normal((T, N)), 0, 1).astype(np.float32)
    p2[:, ::501] = p1[::501]
    p2[1, 70_000:80_500] = p1[70_000:80_500] + np.float32(0.25)  # a plateau of ties across a shard boundary
    y = (rng.random(N) < 0.5).astype(np.float32)
    ridx, ractive = oracle.disparity_select(p2, p1, 0.25, 0.05, blind, y)
    edges = [shard_range(N, r, shards) for r in range(shards)]
    # device-resident shards (what a rank holds after scoring)
    dp2 = [torch.from_numpy(np.ascontiguousarray(p2[:, lo:hi])).cuda() for lo, hi in edges]
    dp1 = [torch.from_numpy(p1[lo:hi].copy()).cuda() for lo, hi in edges]
    dy = [torch.from_numpy(y[lo:hi].copy()).cuda() for lo, hi in edges]
    m = int(N * 0.25)
    thr = np.zeros((T,), np.uint64)
    if not blind:
        for t in range(T):
            hist = lambda pre, ps: sum(ctx.disparity_hist(dp2[r][t], dp1[r], edges[r][0], pre, ps).astype(np.int64)
                                       for r in range(shards))
            thr[t] = np.uint64(global_threshold_key(hist, m))
    parts, active = [], 0
    for r in range(shards):
        idx, a = ctx.disparity_select_sharded(dp2[r], dp1[r], edges[r][0], thr, 0.05, blind, dy[r])
        parts.append(idx)
        active += a
    np.testing.assert_array_equal(np.concatenate(parts), ridx)
    assert active == ractive


def test_allpairs_vote_entropy_refine_vs_exact_and_oracle(ctx, oracle):
    """Vote entropy (Committee's default strategy) over a probe x gallery block with implicit pair indices, through the
    tensor-core vote filter: identical to the exact mode and to the oracle."""
    pool = synth.make_pool(1, N=64, M=1200)
    P, G, B = 200, 800, 300
    probe = np.ascontiguousarray(pool.table[:P])
    gallery = np.ascontiguousarray(pool.table[250:250 + G])
    ctx.set_table(3, probe)
    ctx.set_table(4, gallery)
    ctx.set_heads(2, pool.heads, compute=_ffi.FP32_EXACT)
    ctx.set_heads(3, pool.heads, compute=_ffi.FP32_REFINE)
    exact = ctx.query_allpairs(2, 3, 4, 0, P, "vote_entropy", B)
    refined = ctx.query_allpairs(3, 3, 4, 0, P, "vote_entropy", B)
    rep = ctx.refine_report()
    assert rep["version"] == 2 and rep["fallback"] == 0, rep
    for a, b in zip(refined, exact):
        np.testing.assert_array_equal(a.view(np.uint64) if a.dtype == np.uint64 else a, b)
    ia, ib = _allpairs_indices(P, G)
    ref = oracle.score(probe, gallery, ia, ib, pool.heads, measure="vote_entropy")
    np.testing.assert_array_equal(refined[2], oracle.topk(ref["util"], B)[2])
