import sys
sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/sq-learn_b200")
import numpy as np
from scipy import stats
import gpu_util as G
from oracle import quantum_utility as qu
from sqlearn import _cabi
z = np.load("/root/repo/tools/_d5_ref.npz")
for name in "abc":
    v = z[name + "_v"]; d = len(v)
    vn = v / np.linalg.norm(v)
    sched = qu.schedule(qu.tomography_N(d, 0.3)).astype(float)
    for early, cache in ((1, 256), (0, 0)):
        _cabi.lib.sq_tomography_tune(7, float(early)); _cabi.lib.sq_tomography_tune(6, float(cache))
        est, n_used, err = G.tomography(np.repeat(v[None], 20000, axis=0), 0.3, sched, stop=1, seed=3)
        print(name, "early", early, "err mean gpu %.5f ref %.5f oracle %.5f | N gpu %.2f ref %.2f | KS err vs ref %.3g vs oracle %.3g | KS N vs ref %.3g" % (
            err.mean(), z[name+"_err"].mean(), z[name+"_oerr"].mean(), n_used.mean(), z[name+"_N"].mean(),
            stats.ks_2samp(err, z[name+"_err"]).pvalue, stats.ks_2samp(err, z[name+"_oerr"]).pvalue, stats.ks_2samp(n_used, z[name+"_N"]).pvalue))
        vals, cnt = np.unique(n_used, return_counts=True); rv, rc = np.unique(z[name+"_N"], return_counts=True)
        print("    N pmf gpu:", dict(zip(vals[:8].astype(int), np.round(cnt[:8]/len(n_used), 4))), " ref:", dict(zip(rv[:8].astype(int), np.round(rc[:8]/len(z[name+"_N"]), 4))))
_cabi.lib.sq_tomography_tune(7, 1.0); _cabi.lib.sq_tomography_tune(6, 256.0)
