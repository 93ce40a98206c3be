"""Generate tests/golden/*.npz from the float64 numpy/scipy transcription of the reference lines
(oracle/ref_numpy.py). The reference's Keras/TF runtime cannot be imported in this image
(SURVEY.md §8c), so these vectors are the committed stand-in: inputs come from alink_b200.synth
(seeded), outputs from ref_numpy in float64. Run from the repo root:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from alink_b200 import synth  # noqa: E402
from oracle import ref_numpy as R  # noqa: E402

CASES = {
    "softmax_d512_k3": dict(cfg=1, N=64, K=3, D=512, act="softmax", kind="arcface", M=1000),
    "sigmoid_d512_k5": dict(cfg=1, N=64, K=5, D=512, act="sigmoid", kind="arcface", M=1000),
    "softmax_d2048_k5": dict(cfg=3, N=64, K=5, D=2048, act="softmax", kind="resnet50", M=512),
}

if __name__ == "__main__":
    for name, kw in CASES.items():
        pool = synth.make_pool(kw["cfg"], N=kw["N"], K=kw["K"], D=kw["D"], act=kw["act"], kind=kw["kind"], M=kw["M"])
        XL, XR = pool.table[pool.ia], pool.table[pool.ib]
        preds = [R.head_forward(XL, XR, *pool.heads.learner(k), act=kw["act"], dtype=np.float64) for k in range(kw["K"])]
        out = dict(per_learner=np.stack(preds, axis=1), consensus=R.bagging_predict(preds),
                   table_sum=np.float64(pool.table.astype(np.float64).sum()), ia=pool.ia)
        for m in R.MEASURES:
            out["util_" + m] = np.asarray(R.utility(m, preds), dtype=np.float64)
        out.update({k: np.array(v) for k, v in kw.items()})
        np.savez_compressed(os.path.join(ROOT, "tests", "golden", name + ".npz"), **out)
        print(name, out["per_learner"].shape, float(out["per_learner"].min()), float(out["per_learner"].max()))
