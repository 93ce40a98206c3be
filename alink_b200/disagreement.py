"""Committee disagreement measures and samplers — modAL.disagreement, which the reference imports at
/root/reference/code/learners.py:12 (`vote_entropy_sampling` is Committee's default strategy, :287).
The source is not in the reference tree; semantics follow SURVEY.md Appendix B:

    vote_entropy        H(vote counts / K)            (Counter loop upstream; fused on the GPU here)
    consensus_entropy   H(mean_k p_k)
    KL_max_disagreement max_k KL(p_k || mean_k p_k)

All three run inside the same fused call as the scoring (K-learner vote/entropy in the epilogue).
"""
from __future__ import annotations

from . import _ffi, engine
from .uncertainty import _n_rows, _selected_rows, shuffled_argmax


def _committee_specs(committee):
    specs = engine.specs_of(committee)
    if specs is None:
        raise TypeError("committee disagreement needs alink_b200 learners (no host fallback)")
    return specs


def _measure(committee, X, measure):
    ctx = X.ctx if isinstance(X, engine.PairPool) else _ffi.default_context()
    engine.score(_committee_specs(committee), X, want_per=False, want_cons=False, ctx=ctx)
    return ctx.utility(measure, _n_rows(X))


def vote_entropy(committee, X, **predict_proba_kwargs):
    return _measure(committee, X, "vote_entropy")


def consensus_entropy(committee, X, **predict_proba_kwargs):
    return _measure(committee, X, "consensus_entropy")


def KL_max_disagreement(committee, X, **predict_proba_kwargs):
    return _measure(committee, X, "kl_max")


def _sample(committee, X, measure, n_instances, random_tie_break):
    if random_tie_break:
        idx = shuffled_argmax(_measure(committee, X, measure), n_instances=n_instances)
    else:
        idx, _, _ = engine.fused_query(_committee_specs(committee), X, measure, n_instances)
    return idx, _selected_rows(X, idx)


def vote_entropy_sampling(committee, X, n_instances=1, random_tie_break=False, **kw):
    return _sample(committee, X, "vote_entropy", n_instances, random_tie_break)


def consensus_entropy_sampling(committee, X, n_instances=1, random_tie_break=False, **kw):
    return _sample(committee, X, "consensus_entropy", n_instances, random_tie_break)


def max_disagreement_sampling(committee, X, n_instances=1, random_tie_break=False, **kw):
    return _sample(committee, X, "kl_max", n_instances, random_tie_break)
