"""Golden vectors for QLSSVC, made by running the reference's own class (sklearn/svm/_qSVM.py, loaded unmodified
by oracle/refload.load_qsvm).  Only runs in the build container (needs /root/reference):

    python tests/golden/gen_golden_qsvm.py        ->  tests/golden/qsvm_reference.npz
"""
import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import refload  # noqa: E402

util = refload.load_utility()
QLSSVC = refload.load_qsvm(util)
rs = np.random.RandomState(77)
N, NT, d = 80, 30, 6
X = rs.randn(N, d) * np.array([1.0, 2.0, 0.5, 1.5, 1.0, 0.7])
w = rs.randn(d)
y = np.where(X @ w + 0.3 * rs.randn(N) > 0, 1.0, -1.0)
Xt = rs.randn(NT, d) * 1.2
out = dict(X=X, y=y, Xt=Xt)

CONFIGS = {
    "lin": dict(kernel='linear', penalty=2.0),
    "rbf": dict(kernel='rbf', penalty=1.0),
    "poly": dict(kernel='poly', penalty=0.5, degree=3, coef0=1.0),
    "sig": dict(kernel='sigmoid', penalty=1.5, coef0=0.5, gamma='auto'),
    "linlr": dict(kernel='linear', penalty=2.0, low_rank=True, var=0.9),
    "rbflr": dict(kernel='rbf', penalty=1.0, low_rank=True, var=5),
    "rbfg": dict(kernel='rbf', penalty=1.0, gamma=0.3),      # a float gamma is ignored by the reference (:392-404)
}
with warnings.catch_warnings():
    warnings.simplefilter("ignore")
    for name, kw in CONFIGS.items():
        m = QLSSVC(**kw).fit(X, y)
        out[name + "_b"] = m.b
        out[name + "_alpha"] = m.alpha
        out[name + "_cond"] = m.cond
        out[name + "_normF"] = m.normF
        out[name + "_sv"] = m.singular_values_F_
        out[name + "_Nu"] = m.Nu
        out[name + "_alpha_F"] = m.alpha_F
        if kw["kernel"] == "linear":
            out[name + "_coef"] = m.coef_
        out[name + "_h"] = m.get_h(Xt)
        out[name + "_betas"] = m.get_betas(Xt)
        out[name + "_P"] = m.get_P(Xt)
        out[name + "_cpred"] = m.classical_predict(Xt)
        out[name + "_train_cx"] = m.get_training_complexity()
        out[name + "_class_cx"] = m.get_classification_complexity(Xt)
        out[name + "_class_cx_rel"] = m.get_classification_complexity(Xt, relative_error=True)
        att = m.get_all_attributes(Xt)
        out[name + "_rel_comp"], out[name + "_abs_comp"] = att[4], att[5]
    # noisy paths: repeated predictions (per-sample frequency of +1) and noisy P / h of one sample
    np.random.seed(5)
    for name, kw in {"absn": dict(kernel='rbf', penalty=1.0, absolute_error=40.0),
                     "reln": dict(kernel='rbf', penalty=1.0, error_type='relative', relative_error=0.9)}.items():
        m = QLSSVC(**kw).fit(X, y)
        reps = 300
        out[name + "_pred"] = np.stack([m.predict(Xt) for _ in range(reps)])
        out[name + "_Papprox"] = np.stack([m.get_P(Xt[:4], approx=True) for _ in range(reps)])
        out[name + "_happrox"] = np.stack([m.get_h(Xt[:4], approx=True) for _ in range(reps)])
        out[name + "_P"] = m.get_P(Xt)
        out[name + "_betas"] = m.get_betas(Xt)
        out[name + "_h"] = m.get_h(Xt)
    m = QLSSVC(kernel='linear', penalty=2.0, relative_error=0.8).fit(X, y)
    hp = [m.get_approximated_hyperplane(Xt[:1]) for _ in range(300)]
    out["hyp_b"] = np.array([float(h[0]) for h in hp])
    out["hyp_coef"] = np.stack([h[1] for h in hp])
    out["hyp_beta"] = m.get_betas(Xt[:1])
    rr = [m.relative_error_routine(0.1, 7.0, 1.3) for _ in range(2000)]
    out["rer"] = np.array(rr)
np.savez_compressed(os.path.join(HERE, "qsvm_reference.npz"), **out)
print("written", os.path.join(HERE, "qsvm_reference.npz"), {k: np.shape(v) for k, v in list(out.items())[:8]})
