import sys, os
sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/sq-learn_b200")
import numpy as np
from scipy import stats
import gpu_util as G
from oracle import quantum_utility as qu
ref = np.load("/root/repo/tests/golden/stochastic_reference_r2.npz")
prob = np.load("/root/repo/tests/golden/r2_problems.npz")
TOMO = {"tomo16_default": ("v16", 0.3, "L2", True, None), "tomo128_default": ("v128", 0.25, "L2", True, None),
 "tomo128_default_inf": ("v128", 0.25, "inf", True, None), "tomo128_single_L2": ("v128", 0.25, "L2", False, None),
 "tomo128_single_inf": ("v128", 0.25, "inf", False, None), "tomo784_default": ("v784", 0.5, "L2", True, None),
 "tomo784_default_inf": ("v784", 0.1, "inf", True, None), "tomo784_single_inf": ("v784", 0.1, "inf", False, None),
 "tomo784_single_L2": ("v784", 0.5, "L2", False, None), "tomo16_single_inf": ("v16", 0.3, "inf", False, None)}
rng = np.random.default_rng(5)
for name,(vname, delta, norm, inc, N) in TOMO.items():
    v = prob[vname]; d = len(v)
    Nfull = qu.tomography_N(d, delta, norm)
    sched = qu.schedule(Nfull).astype(float) if inc else [float(Nfull)]
    for seed in (1, 2):
        est, n_used, err = G.tomography(np.repeat(v[None], 10000, axis=0), delta, sched, norm_inf=int(norm=="inf"), stop=int(inc), seed=seed)
        r_err, r_N = ref[name+"__err"], ref[name+"__N"]
        print(name, "seed", seed, "KS err p=%.3g  mean err mine %.5f ref %.5f | N mine %.1f ref %.1f | KS c0 p=%.3g" % (
            stats.ks_2samp(err, r_err).pvalue, err.mean(), r_err.mean(), n_used.mean(), r_N.mean(), stats.ks_2samp(est[:,0], ref[name+"__c0"]).pvalue), flush=True)
    # oracle (numpy multinomial) for comparison
    if d <= 128 or not inc:
        o = []
        for _ in range(600 if d > 128 else 3000):
            r = qu.real_tomography(v, delta=delta, norm=norm, incremental_measure=inc, rng=rng)
            k = list(r.keys())[-1]
            o.append((np.linalg.norm(v - r[k], ord=2 if norm=="L2" else np.inf), k))
        o = np.array(o)
        print("   oracle vs ref: KS err p=%.3g mean %.5f N %.1f ; oracle vs mine p=%.3g" % (stats.ks_2samp(o[:,0], r_err).pvalue, o[:,0].mean(), o[:,1].mean(), stats.ks_2samp(o[:,0], err).pvalue), flush=True)
