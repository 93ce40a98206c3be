"""Golden vectors produced by EXECUTING THE REFERENCE'S OWN CODE (VERDICT r1 "next" item 1).

Run here (the container that has /root/reference), from the repo root:
    python tests/golden/make_ref_fixtures.py
It writes tests/golden/ref_*.npz. The tests (tests/test_ref_pinned.py, tests/test_gpu_ref_pinned.py) only read
the committed .npz files: /root/reference does not exist on the GPU box.

What runs UNMODIFIED from /root/reference/code (imported as modules, or exec'd as source text):
    uncertainty.py    _proba_uncertainty :15-26, _proba_margin :29-46, _proba_entropy :49-60,
                      classifier_uncertainty/_margin/_entropy :63-130, uncertainty/margin/entropy_sampling :133-217
    committee.py      Bagging.predict :13-20
    learners.py       ActiveLearner, Committee.vote / vote_proba / predict_proba / predict :316-416 (on base.py)
    keras_wrapper.py  KerasClassifier.predict :235-242, predict_proba :263-270
    readDFW.py        createMiniBatch :222-244          readMTP.py  createMiniBatch :123-135
    ALINK.py          the disparity / grey-zone / oracle-agreement block :170-201 (exec of those source lines)
What is STUBBED, because the package is not installable in this image (no network): `modAL.*`, `keras.*`, `noise`.
The stubs supply only what the imports need. The two modAL functions that take part in the arithmetic are restated
from upstream modAL: multi_argmax = np.argpartition(-values, n-1)[:n]; shuffled_argmax is not
exercised. Everything else on the pinned rows — 1 - max, partition margin, scipy.stats.entropy, np.sum / K,
np.mean in float64, argmax, hstack widening, argsort prefix, set intersection, the Python pair loops — is the
reference's own line, executed by this interpreter with the installed numpy / scipy.
NOT pinned here: S1-S3, the Dense / softmax / sigmoid graph (code/siamese.py:24-33), which only exists inside
Keras / TensorFlow. Its inputs to the rows above — per-learner probabilities — come from the C oracle.
"""
import hashlib
import importlib
import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = "/root/reference/code"
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "tests", "golden")


def install_stubs():
    """sys.modules entries for the packages the reference imports but this image lacks."""
    def mod(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    def multi_argmax(values, n_instances=1):  # modAL.utils.selection [external, SURVEY App. B]
        assert n_instances <= len(values)
        return np.argpartition(-values, n_instances - 1, axis=0)[:n_instances]

    def shuffled_argmax(values, n_instances=1):
        perm = np.random.permutation(len(values))
        order = np.argsort(values[perm], kind="mergesort")[len(values) - n_instances:]
        return perm[order]

    def check_class_labels(*estimators):  # modAL.utils.validation [external]
        cls = [e.classes_ for e in estimators]
        return all(np.array_equal(cls[0], c) for c in cls[1:])

    def check_class_proba(proba, known_labels, all_labels):
        out = np.zeros((proba.shape[0], len(all_labels)))
        for i, lab in enumerate(known_labels):
            out[:, list(all_labels).index(lab)] = proba[:, i]
        return out

    def data_vstack(blocks):
        return np.concatenate(blocks)

    def _unused(*a, **k):
        raise NotImplementedError("not on the pinned path")

    mod("modAL")
    mod("modAL.utils")
    mod("modAL.utils.data", modALinput=object, data_vstack=data_vstack)
    mod("modAL.utils.selection", multi_argmax=multi_argmax, shuffled_argmax=shuffled_argmax)
    mod("modAL.utils.validation", check_class_labels=check_class_labels, check_class_proba=check_class_proba)
    mod("modAL.uncertainty", uncertainty_sampling=_unused)
    mod("modAL.disagreement", vote_entropy_sampling=_unused, max_std_sampling=_unused)
    mod("modAL.acquisition", max_EI=_unused)
    mod("noise")  # code/committee.py:3 imports it; Bagging.predict does not use it

    class Sequential:  # signatures only: keras_wrapper's has_arg() filters kwargs against them
        def fit(self, x=None, y=None, batch_size=None, epochs=1, verbose=1, **kw): ...
        def predict(self, x, batch_size=None, verbose=0, steps=None): ...
        def predict_classes(self, x, batch_size=32, verbose=0): ...
        def predict_proba(self, x, batch_size=32, verbose=0): ...
        def evaluate(self, x=None, y=None, batch_size=None, verbose=1, **kw): ...

    def has_arg(fn, name, accept_all=False):
        import inspect
        return name in inspect.signature(fn).parameters

    mod("keras")
    mod("keras.utils")
    mod("keras.utils.np_utils", to_categorical=lambda y, n=None: np.eye(n or int(np.max(y)) + 1)[np.asarray(y, int)])
    mod("keras.utils.generic_utils", has_arg=has_arg)
    mod("keras.models", Sequential=Sequential)


def import_reference():
    """The reference modules, unmodified, under their own names."""
    install_stubs()
    sys.path.insert(0, REF)
    try:
        names = ("uncertainty", "committee", "base", "learners", "keras_wrapper", "readDFW", "readMTP")
        return {n: importlib.import_module(n) for n in names}
    finally:
        sys.path.remove(REF)


def alink_block_source():
    """Source text of code/ALINK.py:170-201 (the disparity rule, grey zone and oracle-agreement filter), dedented."""
    with open(os.path.join(REF, "ALINK.py")) as f:
        lines = f.read().split("\n")
    blk = lines[170:201]  # 1-based lines 171-201 (170 is the disguised-model predict call that feeds it)
    assert blk[0].strip() == "misclassifiedIndices = []", blk[0]
    assert blk[-1].strip() == "queryIndices.append(j)", blk[-1]
    return "\n".join(l[2:] if l.startswith("\t\t") else l for l in blk)


class FakeNet:
    """Stands where SiameseNetwork stands: `predict` returns this learner's (N, C) fp32 probabilities."""
    loss = "categorical_crossentropy"

    def __init__(self, probs):
        self.probs = probs

    def predict(self, x, **kw):
        return self.probs


class FakeClassifier:
    """An estimator whose predict_proba returns given class probabilities (what uncertainty.py consumes)."""
    def __init__(self, proba):
        self.proba = proba

    def predict_proba(self, X, **kw):
        return self.proba


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def case_inputs(name):
    """Per-learner probabilities of the C oracle on a seeded synthetic pool (the tests regenerate the same)."""
    from alink_b200 import synth
    from oracle import oracle_ffi
    oracle_ffi.build()
    if name == "cfg1":       # BASELINE configs[0]: 2 000 pairs, 512-d, K = 3, top-100
        pool = synth.make_pool(1)
    elif name == "cfg2_50k":  # a 50 k-pair slice of configs[1]: K = 5, top-1000
        pool = synth.make_pool(2, N=50_000)
    elif name == "cfg1_sigmoid":  # siamese3-style head: one sigmoid column
        pool = synth.make_pool(1, act="sigmoid")
    else:
        raise KeyError(name)
    out = oracle_ffi.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads, want_util=False)
    return pool, out["per_learner"]


def run_measures(ref, per, B):
    """Rows B1, U1-U3, Q1 on per-learner probabilities per (N, K, 2) fp32."""
    U = ref["uncertainty"]
    N, K, _ = per.shape
    bag = ref["committee"].Bagging([FakeNet(np.ascontiguousarray(per[:, k, :])) for k in range(K)], [])
    X = [np.zeros((N, 1), np.float32), np.zeros((N, 1), np.float32)]
    cons = bag.predict(X)                                    # committee.py:13-20
    clf = FakeClassifier(cons)
    res = {"consensus": cons,
           "proba_uncertainty": U._proba_uncertainty(cons),  # uncertainty.py:26
           "proba_margin": U._proba_margin(cons),            # :43-44
           "proba_entropy": U._proba_entropy(cons),          # :60
           "classifier_uncertainty": U.classifier_uncertainty(clf, X),  # :82
           "classifier_margin": U.classifier_margin(clf, X),            # :104-110
           "classifier_entropy": U.classifier_entropy(clf, X)}          # :130
    for nm, fn in (("uncertainty", U.uncertainty_sampling), ("margin", U.margin_sampling), ("entropy", U.entropy_sampling)):
        idx, rows = fn(clf, X, n_instances=B)                # :133-217 -> multi_argmax
        res["sel_" + nm] = np.sort(np.asarray(idx, dtype=np.int64))
    # single learner (existing_al.py: one KerasClassifier behind an ActiveLearner)
    clf0 = FakeClassifier(np.ascontiguousarray(per[:, 0, :]))
    res["single_uncertainty"] = U.classifier_uncertainty(clf0, X)
    res["single_sel_uncertainty"] = np.sort(np.asarray(U.uncertainty_sampling(clf0, X, n_instances=B)[0], dtype=np.int64))
    return res


def tie_free(values, sel, B):
    """True when the B-th and (B+1)-th largest values differ (multi_argmax leaves ties unspecified)."""
    s = np.sort(np.asarray(values))[::-1]
    return bool(s[B - 1] != s[B]) and len(sel) == B


def run_committee(ref, per):
    """Rows C1-C4, W1-W2 through the reference's ActiveLearner / Committee / KerasClassifier classes."""
    L, KW = ref["learners"], ref["keras_wrapper"]
    N, K, C = per.shape
    learners = []
    for k in range(K):
        net = FakeNet(np.ascontiguousarray(per[:, k, :]))
        est = KW.KerasClassifier(build_fn=lambda net=net: net)
        est.classes_ = np.arange(2)
        learners.append(L.ActiveLearner(estimator=est))
    com = L.Committee(learner_list=learners)
    X = np.zeros((N, 1), np.float32)  # learners.py:376,396 read X.shape[0]
    out = {"vote_proba": com.vote_proba(X), "predict_proba": com.predict_proba(X), "predict": com.predict(X),
           "kc_predict_proba": learners[0].estimator.predict_proba(X),   # keras_wrapper.py:263-270
           "kc_predict": np.asarray(learners[0].estimator.predict(X))}   # :235-242
    if C == 2:  # with a 1-column sigmoid head KerasClassifier.predict returns (N, 1) and the reference's own
        out["vote"] = com.vote(X)  # Committee.vote (learners.py:379) cannot broadcast it: a quirk, not pinned
    return out


def run_pairs(ref):
    """Row P1: the reference's pair loops on ragged groups; an 'image' is a 1-element vector holding its id."""
    rng = np.random.Generator(np.random.PCG64(7))
    out = {}
    for tag, sizes_plain, sizes_dig in (("a", [2, 1, 3], [1, 2, 2]), ("b", [1, 1, 1, 1, 2], [3, 1, 1, 2, 1]), ("c", [4], [5])):
        nid = 0
        plain, dig = [], []
        for s in sizes_plain:
            plain.append(np.arange(nid, nid + s, dtype=np.float32).reshape(s, 1))
            nid += s
        for s in sizes_dig:
            dig.append(np.arange(nid, nid + s, dtype=np.float32).reshape(s, 1))
            nid += s
        (XL, XR), Y = ref["readDFW"].createMiniBatch(plain, dig)      # readDFW.py:222-244
        out[f"dfw_{tag}_plain"] = np.array(sizes_plain)
        out[f"dfw_{tag}_dig"] = np.array(sizes_dig)
        out[f"dfw_{tag}_left"] = XL[:, 0].astype(np.int64)
        out[f"dfw_{tag}_right"] = XR[:, 0].astype(np.int64)
        out[f"dfw_{tag}_y"] = Y[:, 0].astype(np.int64)
        (XL, XR), Y = ref["readMTP"].createMiniBatch(plain)           # readMTP.py:123-135
        out[f"mtp_{tag}_left"] = XL[:, 0].astype(np.int64)
        out[f"mtp_{tag}_right"] = XR[:, 0].astype(np.int64)
        out[f"mtp_{tag}_y"] = Y[:, 0].astype(np.int64)
    del rng
    return out


def disparity_inputs(pool, per, T=4, seed=11):
    """p1 = Bagging mean of the committee (fp32, (N, 2)); p2[t] = one learner's scores on a perturbed copy of the
    pool (here: learner 0's logit-space jitter — only the numbers matter to the rule); y = genuine labels."""
    rng = np.random.Generator(np.random.PCG64(seed))
    N = per.shape[0]
    p1 = (np.sum(np.array([per[:, k, :] for k in range(per.shape[1])]), axis=0) / per.shape[1]).astype(np.float32)
    p2 = np.empty((T, N, 2), np.float32)
    for t in range(T):
        z = np.log(per[:, 0, 1].astype(np.float64) / np.maximum(per[:, 0, 0].astype(np.float64), 1e-30))
        z = z + rng.normal(0.0, 0.8, size=N)
        q = (1.0 / (1.0 + np.exp(-z))).astype(np.float32)
        p2[t, :, 1] = q
        p2[t, :, 0] = np.float32(1.0) - q
    y = pool.genuine.astype(np.float32).reshape(N, 1)
    return p2, p1, y


def run_disparity(src, p2, p1, y, ratio, eps, blind):
    """exec of code/ALINK.py:170-201 with the names that block reads."""
    ns = {"np": np, "Set": set, "FLAGS": types.SimpleNamespace(blind_strategy=blind, disparity_ratio=ratio, eps=eps),
          "disguisedPredictions": [p2[t] for t in range(p2.shape[0])], "ensemblePredictions": p1, "batch_y": y,
          "ACTIVE_COUNT": 0}
    exec(compile(src, "ALINK.py:170-201", "exec"), ns)
    return np.sort(np.asarray(ns["queryIndices"], dtype=np.int64)), int(ns["ACTIVE_COUNT"])


def main():
    ref = import_reference()
    src = alink_block_source()
    manifest = []
    for name, B in (("cfg1", 100), ("cfg2_50k", 1000)):
        pool, per = case_inputs(name)
        res = run_measures(ref, per, B)
        com = run_committee(ref, per)
        out = {"input_sha256": np.array(sha(per)), "B": np.array(B), "N": np.array(per.shape[0]), "K": np.array(per.shape[1])}
        for k_, v in res.items():
            assert v.dtype in (np.float32, np.int64), (k_, v.dtype)
        assert com["vote_proba"].dtype == np.float64 and com["predict_proba"].dtype == np.float64
        for nm, key in (("uncertainty", "proba_uncertainty"), ("margin", "proba_margin"), ("entropy", "proba_entropy")):
            vals = res[key] if nm != "margin" else -res[key]
            out["tie_free_" + nm] = np.array(tie_free(vals, res["sel_" + nm], B))
        out["tie_free_single_uncertainty"] = np.array(tie_free(res["single_uncertainty"], res["single_sel_uncertainty"], B))
        if name == "cfg1":
            out["per_learner"] = per
            out.update(res)
            out.update({"com_" + k_: v for k_, v in com.items()})
        else:  # large case: selections in full, vectors as hashes + a strided sample (entropy in full: tolerance row)
            for k_, v in res.items():
                if k_.startswith("sel_") or k_.startswith("single_sel_") or k_ in ("proba_entropy", "classifier_entropy"):
                    out[k_] = v
                else:
                    out[k_ + "_sha256"] = np.array(sha(v))
                    out[k_ + "_every97"] = v[::97]
            for k_, v in com.items():
                out["com_" + k_ + "_sha256"] = np.array(sha(v))
                out["com_" + k_ + "_every97"] = v[::97]
        # disparity block on this pool's numbers
        p2, p1, y = disparity_inputs(pool, per)
        out["disp_input_sha256"] = np.array(sha(p2) + sha(p1) + sha(y))
        for tag, ratio, eps, blind in (("r25", 0.25, 0.05, False), ("r50", 0.5, 0.1, False), ("blind", 0.25, 0.05, True)):
            q, active = run_disparity(src, p2, p1, y, ratio, eps, blind)
            out[f"disp_{tag}_idx"] = q
            out[f"disp_{tag}_active"] = np.array(active)
            out[f"disp_{tag}_params"] = np.array([ratio, eps, float(blind)])
        np.savez_compressed(os.path.join(OUT, f"ref_{name}.npz"), **out)
        manifest.append((name, per.shape, {k_: (int(out[k_]) if out[k_].dtype == bool else None) for k_ in out if k_.startswith("tie_free")},
                         [int(out[f"disp_{t}_idx"].size) for t in ("r25", "r50", "blind")]))
    # sigmoid-1 head: the widening of keras_wrapper.py:263-270 and predict's `> 0.5`
    pool, per = case_inputs("cfg1_sigmoid")
    com = run_committee(ref, per)
    np.savez_compressed(os.path.join(OUT, "ref_cfg1_sigmoid.npz"), input_sha256=np.array(sha(per)), per_learner=per,
                        **{"com_" + k_: v for k_, v in com.items()})
    manifest.append(("cfg1_sigmoid", per.shape, com["kc_predict_proba"].shape, com["kc_predict"].shape))
    np.savez_compressed(os.path.join(OUT, "ref_pairs.npz"), **run_pairs(ref))
    for m in manifest:
        print(m)


if __name__ == "__main__":
    main()
