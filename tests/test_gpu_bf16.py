"""GPU parity of the BF16_TC (tcgen05) head against the fp32 oracle: |dp| <= 2e-2 (north_star
tolerance for bf16 mode), selected-set overlap reported, fused-epilogue utility == utility kernel."""
import numpy as np
import pytest

from alink_b200 import _ffi, synth

pytestmark = pytest.mark.gpu

TOL = 2e-2  # BASELINE.json north_star: "2e-2 absolute in bf16 mode"


def _install(ctx, pool, group=1, feat_mode=0):
    ctx.set_heads(group, pool.heads, feat_mode=feat_mode, compute=_ffi.BF16_TC)
    ctx.set_table(0, pool.table)


@pytest.mark.parametrize("act", ["softmax", "sigmoid"])
@pytest.mark.parametrize("N", [1, 127, 128, 129, 2000, 40000])
def test_bf16_scores_within_tolerance(ctx, oracle, act, N):
    pool = synth.make_pool(1, N=max(N, 64), act=act)
    ia, ib = pool.ia[:N].copy(), pool.ib[:N].copy()
    _install(ctx, pool)
    per, cons = ctx.score_indexed(1, 0, 0, ia, ib, N, pool.heads.K, pool.heads.C)
    ref = oracle.score(pool.table, pool.table, ia, ib, pool.heads)
    err = np.abs(per - ref["per_learner"]).max()
    assert err <= TOL, f"max |dp| = {err}"
    assert np.abs(cons - ref["consensus"]).max() <= TOL
    assert per.std() > 0.05  # not degenerate


def test_bf16_k5_committee_and_overlap(ctx, oracle):
    pool = synth.make_pool(2, N=50000, M=20000)
    _install(ctx, pool)
    idx, util, keys = ctx.query(1, 0, 0, pool.ia, pool.ib, pool.N, "uncertainty", pool.B)
    ref = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads, measure="uncertainty")
    ridx, rutil, _ = oracle.topk(ref["util"], pool.B)
    overlap = len(set(idx.tolist()) & set(ridx.tolist())) / pool.B
    print(f"bf16 top-{pool.B} overlap with the fp32 oracle: {overlap:.3f}")
    assert overlap > 0.5
    # the utilities the kernel's fused epilogue reports are within tolerance of the oracle's at those pairs
    assert np.abs(util - ref["util"][idx]).max() <= TOL


@pytest.mark.parametrize("measure", ["uncertainty", "neg_margin", "entropy", "vote_entropy", "consensus_entropy", "kl_max"])
def test_bf16_fused_utility_equals_utility_kernel(ctx, measure):
    """The epilogue's K-learner reduction must be bit-identical to utility.cuh on the same probabilities."""
    pool = synth.make_pool(1, N=3000)
    _install(ctx, pool)
    idx, util, keys = ctx.query(1, 0, 0, pool.ia, pool.ib, pool.N, measure, pool.B)
    per, _ = ctx.score_indexed(1, 0, 0, pool.ia, pool.ib, pool.N, pool.heads.K, pool.heads.C)
    u2 = ctx.utility_from_proba(per, measure)
    idx2, util2, keys2 = ctx.topk_values(u2, pool.B)
    np.testing.assert_array_equal(keys, keys2)


def test_bf16_table_dtype_bf16(ctx, oracle):
    import torch
    pool = synth.make_pool(1, N=1000)
    ctx.set_heads(1, pool.heads, compute=_ffi.BF16_TC)
    tb = torch.from_numpy(pool.table).cuda().to(torch.bfloat16).contiguous()
    ctx.set_table(2, tb)
    per, _ = ctx.score_indexed(1, 2, 2, pool.ia, pool.ib, pool.N, pool.heads.K, pool.heads.C)
    ref = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads)
    assert np.abs(per - ref["per_learner"]).max() <= TOL


def test_bf16_smallres_widths(ctx, oracle):
    pool = synth.make_pool(1, N=700, H1=128, H2=32, K=2)
    _install(ctx, pool)
    per, _ = ctx.score_indexed(1, 0, 0, pool.ia, pool.ib, pool.N, pool.heads.K, pool.heads.C)
    ref = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads)
    assert np.abs(per - ref["per_learner"]).max() <= TOL


@pytest.mark.parametrize("act", ["softmax", "sigmoid"])
@pytest.mark.parametrize("K,H1,H2", [(1, 128, 32), (3, 128, 64), (1, 512, 64), (3, 512, 32), (5, 256, 64)])
def test_bf16_item_loop_shapes(ctx, oracle, K, H1, H2, act):
    """Odd committee sizes, one-chunk learners (H1 = 128: every item ends a learner) and several tiles per CTA: the
    layer-2 accumulator and the learner parameters alternate between two buffers by RUNNING learner count, and
    a learner is finished one item late (csrc/head_tc.cuh epilogue) — every (tile, learner) must still land in
    its own row of the output."""
    pool = synth.make_pool(1, N=60000, K=K, H1=H1, H2=H2, act=act, M=3000)
    _install(ctx, pool)
    per, cons = ctx.score_indexed(1, 0, 0, pool.ia, pool.ib, pool.N, pool.heads.K, pool.heads.C)
    ref = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads)
    assert np.abs(per - ref["per_learner"]).max() <= TOL
    assert np.abs(cons - ref["consensus"]).max() <= TOL


def test_bf16_hadamard_d256(ctx, oracle):
    pool = synth.make_pool(1, N=500, D=256, feat_mode=1)
    _install(ctx, pool, feat_mode=1)
    per, _ = ctx.score_indexed(1, 0, 0, pool.ia, pool.ib, pool.N, pool.heads.K, pool.heads.C)
    ref = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads, feat_mode=1)
    assert np.abs(per - ref["per_learner"]).max() <= TOL


# ---- FP32_REFINE: tensor-core filter + bit-exact re-score of the survivors (csrc/refine.cuh)
@pytest.mark.parametrize("measure", ["uncertainty", "neg_margin", "entropy", "consensus_entropy", "kl_max", "vote_entropy"])
def test_refine_selection_is_bit_identical_to_fp32(ctx, oracle, measure):
    pool = synth.make_pool(2, N=60000, M=20000)
    ctx.set_heads(2, pool.heads, compute=_ffi.FP32_REFINE)
    ctx.set_table(0, pool.table)
    idx, util, keys = ctx.query(2, 0, 0, pool.ia, pool.ib, pool.N, measure, pool.B, base=5)
    survivors, max_err = ctx.refine_stats()
    ref = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads, measure=measure)
    ridx, rutil, rkeys = oracle.topk(ref["util"], pool.B, base=5)
    np.testing.assert_array_equal(keys, rkeys)
    np.testing.assert_array_equal(idx, ridx)
    np.testing.assert_array_equal(util.view(np.uint32), rutil.view(np.uint32))
    if measure == "vote_entropy":
        # the vote filter (refine.cuh): only pairs whose votes stayed uncertain after the split stage are re-scored exactly
        rep = ctx.refine_report()
        assert rep["version"] == 2 and rep["fallback"] == 0 and rep["rounds"] == 1, rep
        assert 0 <= survivors < 0.01 * pool.N and 0 < rep["n_stage2"] < 0.25 * pool.N
        assert rep["err1_max"] <= rep["eps1_used"] and rep["err2_max"] <= rep["eps2_used"]
    else:
        assert pool.B <= survivors < pool.N
        scale = {"uncertainty": 1.0, "neg_margin": 2.0, "entropy": 1.25, "consensus_entropy": 1.25, "kl_max": 5.0}[measure]
        d = ctx.refine_detail()
        assert d["filter"] == "f16"  # unit-norm embeddings cannot overflow half: the narrow-band filter runs
        assert d["rounds"] == 1, f"the default bound had to be widened: {d}, max err {max_err}"
        assert max_err <= d["eps_used"] * scale, f"observed filter utility error {max_err} exceeds the bound {d}"
        print(f"refine[{measure}]: {survivors} of {pool.N} pairs re-scored exactly, max err {max_err:.2e}, {d}")


def test_refine_falls_back_to_bf16_when_half_could_overflow(ctx, oracle):
    """Features beyond the half range (|a| + |b| >= 6e4 after scaling the table by 1e6): the filter must use the bf16 images; selection stays exact."""
    pool = synth.make_pool(2, N=30000, M=5000)
    table = (pool.table * np.float32(1.0e6)).astype(np.float32)
    import copy
    heads = copy.deepcopy(pool.heads)
    heads.W1 = [np.ascontiguousarray(w * np.float32(1.0 / 1.0e6), dtype=np.float32) for w in heads.W1]
    ctx.set_heads(2, heads, compute=_ffi.FP32_REFINE)
    ctx.set_table(0, table)
    idx, util, keys = ctx.query(2, 0, 0, pool.ia, pool.ib, pool.N, "uncertainty", pool.B)
    assert ctx.refine_detail()["filter"] == "bf16"
    ref = oracle.score(table, table, pool.ia, pool.ib, heads, measure="uncertainty")
    ridx, rutil, rkeys = oracle.topk(ref["util"], pool.B)
    np.testing.assert_array_equal(keys, rkeys)
    np.testing.assert_array_equal(util.view(np.uint32), rutil.view(np.uint32))


def test_refine_widens_a_bound_that_is_too_tight(ctx, oracle):
    """A caller-supplied eps far below the real filter error is detected on the candidates and widened."""
    pool = synth.make_pool(2, N=60000, M=20000)
    ctx.set_heads(2, pool.heads, compute=_ffi.FP32_REFINE)
    ctx.set_table(0, pool.table)
    ctx.set_refine_eps(1e-6)
    try:
        idx, util, keys = ctx.query(2, 0, 0, pool.ia, pool.ib, pool.N, "uncertainty", pool.B)
        d = ctx.refine_detail()
    finally:
        ctx.set_refine_eps(0.0)
    assert d["rounds"] > 1 and d["eps_used"] > 1e-6
    ref = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads, measure="uncertainty")
    _, _, rkeys = oracle.topk(ref["util"], pool.B)
    np.testing.assert_array_equal(keys, rkeys)


def test_refine_small_pool_and_scores_are_exact(ctx, oracle):
    pool = synth.make_pool(1)  # 2,000 pairs: below the filter's minimum, runs the exact path
    ctx.set_heads(2, pool.heads, compute=_ffi.FP32_REFINE)
    ctx.set_table(0, pool.table)
    idx, util, keys = ctx.query(2, 0, 0, pool.ia, pool.ib, pool.N, "uncertainty", pool.B)
    ref = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads, measure="uncertainty")
    ridx, _, rkeys = oracle.topk(ref["util"], pool.B)
    np.testing.assert_array_equal(keys, rkeys)
    per, cons = ctx.score_indexed(2, 0, 0, pool.ia, pool.ib, pool.N, pool.heads.K, pool.heads.C)
    np.testing.assert_array_equal(per.view(np.uint32), ref["per_learner"].view(np.uint32))  # score calls stay exact


# ---- wide features: the k-outer tensor-core kernel (csrc/head_tc_ko.cuh)
@pytest.mark.parametrize("act", ["softmax", "sigmoid"])
@pytest.mark.parametrize("N", [1, 129, 700, 40000])
def test_bf16_d2048_resnet_pool(ctx, oracle, act, N):
    """2048-d VGGFace-ResNet50-shaped features (code/siamese.py:203-216), K = 2 learners."""
    pool = synth.make_pool(1, N=max(N, 64), D=2048, kind="resnet50", K=2, act=act, M=800)
    ia, ib = pool.ia[:N].copy(), pool.ib[:N].copy()
    _install(ctx, pool)
    per, cons = ctx.score_indexed(1, 0, 0, ia, ib, N, pool.heads.K, pool.heads.C)
    ref = oracle.score(pool.table, pool.table, ia, ib, pool.heads)
    err = np.abs(per - ref["per_learner"]).max()
    assert err <= TOL, f"max |dp| = {err}"
    assert np.abs(cons - ref["consensus"]).max() <= TOL


def test_bf16_d1024_h256_fused_utility(ctx, oracle):
    pool = synth.make_pool(1, N=5000, D=1024, kind="resnet50", K=3, H1=256, H2=32, M=700)
    _install(ctx, pool)
    idx, util, keys = ctx.query(1, 0, 0, pool.ia, pool.ib, pool.N, "entropy", 64)
    per, _ = ctx.score_indexed(1, 0, 0, pool.ia, pool.ib, pool.N, pool.heads.K, pool.heads.C)
    ref = oracle.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads)
    assert np.abs(per - ref["per_learner"]).max() <= TOL
    _, _, keys2 = ctx.topk_values(ctx.utility_from_proba(per, "entropy"), 64)
    np.testing.assert_array_equal(keys, keys2)
