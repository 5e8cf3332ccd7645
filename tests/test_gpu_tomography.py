"""K1 (batched tomography) and its binomial sampler against the reference's
law: fixtures hold samples drawn by the reference's own code; chi-square / KS
at fixed seeds, plus the properties the algorithm guarantees."""
import numpy as np
import pytest
from scipy import stats

pytestmark = pytest.mark.gpu
P_MIN = 1e-3


@pytest.mark.parametrize("n,p", [(1, 0.3), (7, 0.5), (40, 0.01), (40, 0.2), (200, 0.03), (30, 0.9), (1000, 0.5),
                                 (1000, 0.999), (12345, 0.0008), (100000, 0.3), (10 ** 7, 1e-6)])
def test_binomial_sampler_chisquare(cuda, n, p):
    import gpu_util as G
    B = 200000
    out = G.binomial(np.full(B, float(n)), np.full(B, p), np.full(B, 1 - p), seed=n)
    assert out.min() >= 0 and out.max() <= n and np.all(out == np.floor(out))
    lo, hi = int(stats.binom.ppf(1e-7, n, p)), int(stats.binom.ppf(1 - 1e-7, n, p))
    ks = np.arange(lo, hi + 1)
    pmf = stats.binom.pmf(ks, n, p)
    obs = np.array([(out == k).sum() for k in ks], dtype=float) if hi - lo < 400 else \
        np.histogram(out, bins=np.arange(lo - 0.5, hi + 1.5))[0].astype(float)
    # merge sparse cells
    order = np.argsort(pmf)
    exp = pmf * B
    small = exp < 8
    o = np.append(obs[~small], obs[small].sum() + (B - obs.sum()))
    e = np.append(exp[~small], exp[small].sum() + (B - exp.sum()))
    keep = e > 0
    assert stats.chisquare(o[keep], e[keep] * o[keep].sum() / e[keep].sum()).pvalue > P_MIN


def test_binomial_huge_counts_moments(cuda):
    import gpu_util as G
    B = 100000
    for n, p in [(5e12, 0.37), (3e9, 1e-4), (8e14, 0.5)]:
        out = G.binomial(np.full(B, n), np.full(B, p), np.full(B, 1 - p), seed=7)
        z = (out - n * p) / np.sqrt(n * p * (1 - p))
        assert abs(z.mean()) < 5 / np.sqrt(B)
        assert abs(z.std() - 1) < 0.02
        assert stats.kstest(z, "norm").pvalue > P_MIN
    # degenerate inputs
    out = G.binomial(np.array([0.0, 5.0, 5.0]), np.array([0.5, 0.0, 1.0]), np.array([0.5, 1.0, 0.0]))
    assert list(out) == [0.0, 0.0, 5.0]


def _default_runs(G, v, delta, B, seed, norm_inf=0):
    from oracle import quantum_utility as qu
    N = qu.tomography_N(len(v), delta, "inf" if norm_inf else "L2")
    sched = qu.schedule(N).astype(float)
    return G.tomography(np.repeat(v[None], B, axis=0), delta, sched, norm_inf=norm_inf, stop=1, seed=seed)


def test_default_mode_matches_reference_law(cuda, golden_stoch):
    import gpu_util as G
    v = golden_stoch["tomo_v"]
    est, n_used, err = _default_runs(G, v, 0.3, 4000, seed=11)
    assert err.max() <= 0.3
    np.testing.assert_allclose(err, np.linalg.norm(v[None] - est, axis=1), rtol=1e-12)
    assert stats.ks_2samp(err, golden_stoch["tomo_default_err"]).pvalue > P_MIN
    assert stats.ks_2samp(n_used, golden_stoch["tomo_default_N"]).pvalue > P_MIN


@pytest.mark.parametrize("norm", ["L2", "inf"])
def test_single_shot_matches_reference_law(cuda, golden_stoch, norm):
    import gpu_util as G
    from oracle import quantum_utility as qu
    v = golden_stoch["tomo_v"]
    N = qu.tomography_N(16, 0.3, norm)
    est, n_used, err = G.tomography(np.repeat(v[None], 4000, axis=0), 0.3, [N], norm_inf=int(norm == "inf"), seed=12)
    assert (n_used == N).all()
    want = np.linalg.norm(v[None] - est, ord=2 if norm == "L2" else np.inf, axis=1)
    np.testing.assert_allclose(err, want, rtol=1e-12)
    assert stats.ks_2samp(err, golden_stoch[f"tomo_single_{norm}_err"]).pvalue > P_MIN
    assert stats.ks_2samp(est[:, 0], golden_stoch[f"tomo_single_{norm}_c0"]).pvalue > P_MIN
    # theorem: ||est - v|| <= delta except with small probability
    assert (err > 0.3).mean() < 0.01


def test_small_N_sign_errors_and_components(cuda, golden_stoch):
    import gpu_util as G
    v = golden_stoch["tomo_v"]
    B = 6000
    est, _, _ = G.tomography(np.repeat(v[None], B, axis=0), 0.3, [40.0], seed=13)
    wrong = int(np.sum(np.sign(est) != np.sign(v)[None]))
    ref_wrong, ref_total = golden_stoch["tomo_N40_signerr"]
    p = (wrong + ref_wrong) / (B * 16 + ref_total)
    z = (wrong / (B * 16) - ref_wrong / ref_total) / np.sqrt(p * (1 - p) * (1 / (B * 16) + 1 / ref_total))
    assert abs(z) < 4
    assert stats.ks_2samp(est[:, 3], golden_stoch["tomo_N40_c3"]).pvalue > P_MIN
    # squared estimates are multinomial frequencies: they sum to one, multiples of 1/N
    np.testing.assert_allclose((est ** 2).sum(axis=1), 1.0, atol=1e-12)
    np.testing.assert_allclose(est ** 2 * 40, np.round(est ** 2 * 40), atol=1e-9)


def test_stage_one_counts_are_multinomial(cuda):
    """est_i^2 * N ~ Binomial(N, v_i^2) marginally (chi-square on one component)"""
    import gpu_util as G
    rs = np.random.RandomState(5)
    v = rs.uniform(-1, 1, 37)
    v /= np.linalg.norm(v)
    N, B = 60, 20000
    est, _, _ = G.tomography(np.repeat(v[None], B, axis=0), 0.5, [float(N)], seed=14)
    for comp in (0, 17, 36):
        c = np.round(est[:, comp] ** 2 * N).astype(int)
        pmf = stats.binom.pmf(np.arange(N + 1), N, v[comp] ** 2)
        obs = np.bincount(c, minlength=N + 1).astype(float)
        small = pmf * B < 8
        o = np.append(obs[~small], obs[small].sum())
        e = np.append(pmf[~small], pmf[small].sum()) * B
        assert stats.chisquare(o, e * o.sum() / e.sum()).pvalue > P_MIN


def test_normalisation_rule_and_zero_vector(cuda):
    import gpu_util as G
    from oracle import quantum_utility as qu
    rs = np.random.RandomState(6)
    v = rs.uniform(-1, 1, 20)
    v /= np.linalg.norm(v)
    N = qu.tomography_N(20, 0.1)
    V = np.stack([5 * v, 1.005 * v, 0.5 * v, np.zeros(20)])
    est, n_used, err = G.tomography(V, 0.1, [N], seed=15)
    # rows 0 and 2 were renormalised, row 1 (within rtol 1e-2 of unit norm) kept as is (Utility.py:301-304)
    np.testing.assert_allclose(np.linalg.norm(est[:3], axis=1), 1.0, atol=1e-12)
    assert err[0] == pytest.approx(np.linalg.norm(v - est[0]), rel=1e-12)
    assert err[1] == pytest.approx(np.linalg.norm(1.005 * v - est[1]), rel=1e-12)
    assert (est[3] == 0).all() and n_used[3] == 0
    assert (err[:3] < 0.1).all()


@pytest.mark.parametrize("d", [1, 2, 3, 61, 128, 784, 1000, 4096])
def test_dims_and_sharding_invariance(cuda, d):
    """any d up to the fused limit; a vector's result depends on (seed, global id) only"""
    import gpu_util as G
    from oracle import quantum_utility as qu
    rs = np.random.RandomState(d)
    B = 6
    V = rs.uniform(-1, 1, (B, d))
    V /= np.linalg.norm(V, axis=1, keepdims=True)
    delta = 0.25
    N = max(qu.tomography_N(d, delta), 50)
    sched = qu.schedule(N).astype(float)
    est, n_used, err = G.tomography(V, delta, sched, stop=1, seed=21, vec_offset=100)
    assert (err <= delta).all()
    np.testing.assert_allclose(np.linalg.norm(est, axis=1), 1.0, atol=1e-10)
    lo = G.tomography(V[:2], delta, sched, stop=1, seed=21, vec_offset=100)
    hi = G.tomography(V[2:], delta, sched, stop=1, seed=21, vec_offset=102)
    assert np.array_equal(np.vstack([lo[0], hi[0]]), est)
    assert np.array_equal(np.concatenate([lo[1], hi[1]]), n_used)
    other = G.tomography(V, delta, sched, stop=1, seed=22, vec_offset=100)
    if d > 2:
        assert not np.array_equal(other[0], est)


@pytest.mark.parametrize("d,delta,stop", [(128, 0.25, 1), (1000, 0.3, 1), (1000, 0.05, 0), (61, 0.1, 1), (2000, 0.2, 0)])
def test_result_does_not_depend_on_threads_per_vector(cuda, d, delta, stop):
    """the host gives a vector L threads in a small batch and L/8 in a large one; estimates, attempt counts
    and errors must agree bit for bit (every reduction is ordered by L alone, draws are keyed by node / sample)"""
    import gpu_util as G
    from oracle import quantum_utility as qu
    rs = np.random.RandomState(d + stop)
    B = 4200                                     # >= 4096 vectors: the many-vector geometry
    base = rs.uniform(-1, 1, (7, d))
    V = base[rs.randint(0, 7, B)] * rs.uniform(0.5, 2.0, (B, 1))
    N = qu.tomography_N(d, delta)
    sched = qu.schedule(N).astype(float) if stop else np.array([float(N)])
    big = G.tomography(V, delta, sched, stop=stop, seed=33, vec_offset=1000)
    for i in (0, 17, 4199):
        one = G.tomography(V[i:i + 1], delta, sched, stop=stop, seed=33, vec_offset=1000 + i)
        assert np.array_equal(one[0][0], big[0][i])
        assert one[1][0] == big[1][i] and one[2][0] == big[2][i]


def test_trace_and_public_api(cuda, golden_stoch):
    import gpu_util as G
    from oracle import quantum_utility as qu
    from sqlearn import QuantumUtility as Q
    v = golden_stoch["tomo_v"]
    sched = qu.schedule(qu.tomography_N(16, 0.3)).astype(float)
    est, n_used, err, tr, att = G.tomography(v, 0.3, sched, stop=1, seed=31, trace=True)
    a = int(att[0])
    assert n_used[0] == sched[a - 1]
    assert np.array_equal(tr[0, a - 1], est[0])
    errs = np.linalg.norm(v[None] - tr[0, :a], axis=1)
    assert (errs[:-1] > 0.3).all() and errs[-1] <= 0.3
    Q.seed(3)
    res = Q.real_tomography(v, delta=0.3)
    assert list(res.keys()) == [int(s) for s in sched[:len(res)]]
    assert np.linalg.norm(v - np.asarray(list(res.values())[-1])) <= 0.3
    out = Q.tomography(np.stack([v, -v]), 0.3)
    assert out.shape == (2, 16) and (np.linalg.norm(out - np.stack([v, -v]), axis=1) <= 0.3).all()
    assert Q.tomography(v, 0) is v
    with pytest.raises(AssertionError):
        Q.tomography(v, -1)
    g = Q.tomography(v, 0.3, true_tomography=False)
    assert np.abs(g - v).max() <= 0.3 / 4
    single = Q.real_tomography(v, delta=0.3, incremental_measure=False)
    assert list(single.keys()) == [qu.tomography_N(16, 0.3)]


def _big(V, delta, sched, norm_inf=0, stop=0, seed=1, vec_offset=0):
    import torch
    from sqlearn import _cabi, _device as dv
    Vd = torch.as_tensor(np.ascontiguousarray(np.atleast_2d(V))).cuda()
    B, d = Vd.shape
    sched = np.ascontiguousarray(sched, dtype=np.float64)
    est = torch.empty_like(Vd)
    n_used = torch.empty(B, dtype=torch.float64, device="cuda")
    err = torch.empty(B, dtype=torch.float64, device="cuda")
    nb = _cabi.lib.sq_tomography_workspace_bytes(B, d)
    ws = torch.empty(max(nb, 8), dtype=torch.uint8, device="cuda")
    _cabi.check(_cabi.lib.sq_tomography_batched_ws(Vd.data_ptr(), B, d, float(delta), norm_inf, stop, sched.ctypes.data,
                                                   len(sched), seed, vec_offset, est.data_ptr(), n_used.data_ptr(),
                                                   err.data_ptr(), ws.data_ptr(), nb, dv.stream_ptr()), "tomo_ws")
    torch.cuda.synchronize()
    return est.cpu().numpy(), n_used.cpu().numpy(), err.cpu().numpy()


@pytest.mark.parametrize("d", [4097, 5000, 20000, 131072, 1000000])
def test_long_vectors(cuda, d):
    """d beyond the fused path: unit-norm output, the stop rule holds, results do not depend on batching"""
    from oracle import quantum_utility as qu
    rs = np.random.RandomState(d % 1000)
    B = 3 if d <= 131072 else 1
    V = rs.uniform(-1, 1, (B, d))
    V[0] *= 3.0                                      # gets normalised (Utility.py:301-304)
    Vn = V / np.linalg.norm(V, axis=1, keepdims=True)
    delta = 0.3
    N = qu.tomography_N(d, delta)
    est, n_used, err = _big(V, delta, [N], seed=5, vec_offset=10)
    assert (n_used == N).all()
    np.testing.assert_allclose(np.linalg.norm(est, axis=1), 1.0, atol=1e-9)
    np.testing.assert_allclose(err, np.linalg.norm(Vn - est, axis=1), rtol=1e-10)
    assert (err < delta).all() and (err > 0).all()
    # squared estimates are counts / N
    np.testing.assert_allclose((est ** 2).sum(axis=1), 1.0, atol=1e-9)
    sched = qu.schedule(N).astype(float)
    e2, n2, r2 = _big(V, delta, sched, stop=1, seed=6, vec_offset=10)
    assert (r2 <= delta).all() and (n2 < N).all()
    np.testing.assert_allclose(r2, np.linalg.norm(Vn - e2, axis=1), rtol=1e-10)
    if B > 1:
        a = _big(V[:1], delta, sched, stop=1, seed=6, vec_offset=10)
        b = _big(V[1:], delta, sched, stop=1, seed=6, vec_offset=11)
        assert np.array_equal(np.vstack([a[0], b[0]]), e2) and np.array_equal(np.concatenate([a[1], b[1]]), n2)
    # L-infinity
    Ni = qu.tomography_N(d, 0.05, "inf")
    e3, _, r3 = _big(V, 0.05, [Ni], norm_inf=1, seed=7)
    np.testing.assert_allclose(r3, np.abs(Vn - e3).max(axis=1), rtol=1e-10)
    assert (r3 < 0.05).all()


@pytest.mark.parametrize("norm_inf", [0, 1])
def test_sharded_phases_equal_the_single_call(cuda, norm_inf):
    """sq_tomography_sharded_phase with the leaf blocks dealt to 3 emulated ranks (own workspace and output each,
    block masses and error partials merged the way the host's all-reduce does) gives sq_tomography_batched_ws's
    result bit for bit, stop rule included"""
    import ctypes
    import torch
    from oracle import quantum_utility as qu
    from sqlearn import _cabi, _device as dv
    rs = np.random.RandomState(12)
    B, d, delta, R = 2, 50000, 0.3 if not norm_inf else 0.02, 3
    V = rs.uniform(-1, 1, (B, d))
    sched = qu.schedule(qu.tomography_N(d, delta, "inf" if norm_inf else "L2")).astype(float)[-12:]
    want = _big(V, delta, sched, stop=1, norm_inf=norm_inf, seed=21, vec_offset=4)
    Vd = torch.as_tensor(V).cuda()
    lay = (ctypes.c_longlong * 8)()
    _cabi.check(_cabi.lib.sq_tomography_workspace_layout(B, d, lay), "layout")
    nblk, tb, nb = int(lay[0]), int(lay[1]), int(lay[7])
    assert nblk == -(-d // tb) and nb == _cabi.lib.sq_tomography_workspace_bytes(B, d)
    ranks = []
    for r in range(R):
        ws = torch.zeros(nb, dtype=torch.uint8, device="cuda")
        f = lambda i, ws=ws: ws[int(lay[2 + i]):int(lay[2 + i]) + B * nblk * 8].view(torch.float64).view(B, nblk)
        ranks.append(dict(ws=ws, est=torch.zeros_like(Vd), n_used=torch.zeros(B, dtype=torch.float64, device="cuda"),
                          err=torch.zeros(B, dtype=torch.float64, device="cuda"), bm_a=f(0), bm_b=f(1), errp=f(3),
                          done=ws[int(lay[6]):int(lay[6]) + 4 * B].view(torch.int32),
                          lo=nblk * r // R, hi=nblk * (r + 1) // R))

    def phase(rk, ph, at, n_meas):
        _cabi.check(_cabi.lib.sq_tomography_sharded_phase(
            Vd.data_ptr(), B, d, delta, norm_inf, 1, float(n_meas), at, 21, 4, rk["est"].data_ptr(),
            rk["n_used"].data_ptr(), rk["err"].data_ptr(), rk["ws"].data_ptr(), nb, rk["lo"], rk["hi"], ph,
            dv.stream_ptr()), "sharded_phase")

    def merge(key, op):
        for rk in ranks:
            rk[key][:, :rk["lo"]] = 0
            rk[key][:, rk["hi"]:] = 0
        tot = ranks[0][key].clone()
        for rk in ranks[1:]:
            tot = op(tot, rk[key])
        for rk in ranks:
            rk[key].copy_(tot)
    for rk in ranks:
        phase(rk, 0, 0, 1.0)
    for at, n_meas in enumerate(sched):
        for rk in ranks:
            phase(rk, 1, at, n_meas)
        merge("bm_a", torch.add); merge("bm_b", torch.add)
        for rk in ranks:
            phase(rk, 2, at, n_meas)
        merge("errp", torch.maximum if norm_inf else torch.add)
        for rk in ranks:
            phase(rk, 3, at, n_meas)
        if bool(ranks[0]["done"].all()):
            break
    est = torch.zeros_like(Vd)
    for rk in ranks:
        est[:, rk["lo"] * tb:rk["hi"] * tb] = rk["est"][:, rk["lo"] * tb:rk["hi"] * tb]
    torch.cuda.synchronize()
    assert np.array_equal(est.cpu().numpy(), want[0])
    for rk in ranks:
        assert np.array_equal(rk["n_used"].cpu().numpy(), want[1]) and np.array_equal(rk["err"].cpu().numpy(), want[2])
    assert (want[1] < sched[-1]).any() or (want[2] > delta).any()     # the stop rule was exercised (or never met)
    # argument checks
    assert _cabi.lib.sq_tomography_sharded_phase(Vd.data_ptr(), B, d, delta, 0, 1, 10.0, 0, 1, 0, est.data_ptr(),
                                                 ranks[0]["n_used"].data_ptr(), ranks[0]["err"].data_ptr(),
                                                 ranks[0]["ws"].data_ptr(), nb, 0, nblk + 1, 1, dv.stream_ptr()) != 0


def test_long_vector_law_matches_oracle(cuda):
    """d = 5000 at a small N: error and one component's estimate follow the oracle's law"""
    from oracle import quantum_utility as qu
    rs = np.random.RandomState(9)
    d, N, B = 5000, 20000, 400
    v = rs.uniform(-1, 1, d)
    v /= np.linalg.norm(v)
    big = np.argmax(np.abs(v))
    est, _, err = _big(np.repeat(v[None], B, axis=0), 0.5, [float(N)], seed=8)
    rng = np.random.default_rng(1)
    o = np.array([qu.tomography_attempt(v, N, rng) for _ in range(B)])
    oerr = np.linalg.norm(v[None] - o, axis=1)
    assert stats.ks_2samp(err, oerr).pvalue > P_MIN
    assert stats.ks_2samp(est[:, big], o[:, big]).pvalue > P_MIN
    wrong, owrong = np.sum(np.sign(est) != np.sign(v)[None]), np.sum(np.sign(o) != np.sign(v)[None])
    assert abs(wrong - owrong) < 6 * np.sqrt(wrong + owrong)


def test_long_vector_public_api(cuda):
    from sqlearn import QuantumUtility as Q
    rs = np.random.RandomState(3)
    v = rs.uniform(-1, 1, 6000)
    Q.seed(4)
    out = Q.tomography(np.stack([v, -2 * v]), 0.4)
    vn = v / np.linalg.norm(v)
    assert out.shape == (2, 6000)
    assert np.linalg.norm(out[0] - vn) <= 0.4 and np.linalg.norm(out[1] + vn) <= 0.4
    res = Q.real_tomography(v, delta=0.4)
    assert len(res) == 1 and np.linalg.norm(np.asarray(list(res.values())[0]) - vn) <= 0.4


def test_speculative_lanes_equal_the_in_order_search(cuda):
    """the attempt-lane search returns exactly what the sequential loop returns"""
    import torch
    from oracle import quantum_utility as qu
    from sqlearn import _cabi, _device as dv
    import gpu_util as G
    rs = np.random.RandomState(11)
    for B, d, delta in [(64, 128, 0.25), (7, 1000, 0.2), (300, 16, 0.3), (5, 3, 0.01)]:
        V = rs.uniform(-1, 1, (B, d))
        V[0] = 0.0                                    # a zero vector in the batch
        sched = qu.schedule(max(qu.tomography_N(d, delta), 50)).astype(float)
        want = G.tomography(V, delta, sched, stop=1, seed=77, vec_offset=5)
        Vd = torch.as_tensor(V).cuda()
        est = torch.empty_like(Vd)
        n_used = torch.empty(B, dtype=torch.float64, device="cuda")
        err = torch.empty(B, dtype=torch.float64, device="cuda")
        att = torch.zeros(B, dtype=torch.int32, device="cuda")
        scratch = torch.zeros(B, dtype=torch.int32, device="cuda")
        for lanes in (1, 8, 13):
            _cabi.check(_cabi.lib.sq_tomography_batched_spec(
                Vd.data_ptr(), B, d, delta, 0, sched.ctypes.data, len(sched), lanes, 77, 5, est.data_ptr(),
                n_used.data_ptr(), err.data_ptr(), att.data_ptr(), scratch.data_ptr(), dv.stream_ptr()), "spec")
            torch.cuda.synchronize()
            assert np.array_equal(est.cpu().numpy(), want[0])
            assert np.array_equal(n_used.cpu().numpy(), want[1])
            assert np.array_equal(err.cpu().numpy(), want[2])
        # the same in one launch (the lane that finishes last replays); the state is prepared once and every
        # call leaves it prepared: three calls in a row, the second with another seed
        state = torch.empty(2 * B, dtype=torch.int32, device="cuda")
        _cabi.check(_cabi.lib.sq_tomography_spec_state_init(state.data_ptr(), B, dv.stream_ptr()), "state")
        other = G.tomography(V, delta, sched, stop=1, seed=78, vec_offset=5)
        for seed, ref in ((77, want), (78, other), (77, want)):
            est.zero_(); n_used.zero_(); err.zero_()
            _cabi.check(_cabi.lib.sq_tomography_batched_spec_fused(
                Vd.data_ptr(), B, d, delta, 0, sched.ctypes.data, len(sched), 8, seed, 5, est.data_ptr(),
                n_used.data_ptr(), err.data_ptr(), att.data_ptr(), state.data_ptr(), dv.stream_ptr()), "spec_fused")
            torch.cuda.synchronize()
            assert np.array_equal(est.cpu().numpy(), ref[0])
            assert np.array_equal(n_used.cpu().numpy(), ref[1])
            assert np.array_equal(err.cpu().numpy(), ref[2])
        assert (state[:B] == 0x7fffffff).all() and (state[B:] == 0).all()


@pytest.mark.parametrize("B,d,delta,norm_inf", [(700, 1000, 0.1, 0), (5000, 128, 0.25, 0), (400, 4096, 0.2, 0),
                                                (3000, 61, 0.1, 0), (900, 300, 0.02, 1), (600, 2, 0.05, 0),
                                                (5000, 3, 0.02, 0), (1200, 33, 0.1, 1)])
def test_group_kernel_equals_the_fused_kernel(cuda, B, d, delta, norm_inf):
    """a single attempt of many vectors runs in the group kernel (sq_tomography_batched_ws with a workspace):
    same draws, same order of additions, so estimate / N / error equal the per-vector kernel's bit for bit -
    with vectors that are renormalised, a zero vector, a ragged last group and a non-zero stream offset"""
    import gpu_util as G
    from oracle import quantum_utility as qu
    from sqlearn import _cabi
    rs = np.random.RandomState(B + d)
    V = rs.uniform(-1, 1, (B, d))
    V /= np.linalg.norm(V, axis=1, keepdims=True)
    V[1] *= 3.0
    V[2] = 0.0
    V[B - 1] *= 0.2
    N = qu.tomography_N(d, delta, "inf" if norm_inf else "L2")
    assert _cabi.lib.sq_tomography_workspace_bytes(B, d) > 0            # the group path is taken
    got = _big(V, delta, [float(N)], norm_inf=norm_inf, seed=91, vec_offset=12345)
    want = G.tomography(V, delta, [float(N)], norm_inf=norm_inf, seed=91, vec_offset=12345)
    for a, b in zip(got, want):
        assert np.array_equal(a, b)
    assert (got[0][2] == 0).all() and got[1][2] == 0
    # a schedule without the stop rule returns its last attempt, keyed by that attempt's index
    sched = qu.schedule(N).astype(float)[-3:]
    got = _big(V, delta, sched, norm_inf=norm_inf, stop=0, seed=92)
    want = G.tomography(V, delta, sched, norm_inf=norm_inf, stop=0, seed=92)
    for a, b in zip(got, want):
        assert np.array_equal(a, b)


@pytest.mark.parametrize("d,delta,norm_inf", [(128, 0.25, 0), (61, 0.1, 0), (1000, 0.3, 0), (300, 0.1, 1), (16, 0.3, 0)])
def test_skipping_hopeless_attempts_changes_nothing(cuda, d, delta, norm_inf):
    """incremental mode gives up an attempt after stage 1 when no choice of signs could meet delta, and does not
    start attempts whose N leaves too many components at zero; switched off (sq_tomography_tune 7), every
    attempt runs both stages - the returned attempt, its N and its error must be the same bits"""
    import gpu_util as G
    from oracle import quantum_utility as qu
    from sqlearn import _cabi
    rs = np.random.RandomState(d)
    B = 300
    V = rs.uniform(-1, 1, (B, d)) * rs.uniform(0.3, 3.0, (B, 1))
    V[7] = 0.0
    sched = qu.schedule(max(qu.tomography_N(d, delta, "inf" if norm_inf else "L2"), 50)).astype(float)
    try:
        _cabi.check(_cabi.lib.sq_tomography_tune(7, 0.0), "tune")
        _cabi.check(_cabi.lib.sq_tomography_tune(6, 0.0), "tune")
        plain = G.tomography(V, delta, sched, norm_inf=norm_inf, stop=1, seed=55, vec_offset=9)
    finally:
        _cabi.check(_cabi.lib.sq_tomography_tune(7, 1.0), "tune")
        _cabi.check(_cabi.lib.sq_tomography_tune(6, 256.0), "tune")
    fast = G.tomography(V, delta, sched, norm_inf=norm_inf, stop=1, seed=55, vec_offset=9)
    for a, b in zip(fast, plain):
        assert np.array_equal(a, b)
    assert (fast[2] <= delta).all()
    # the trace of every attempt is still complete (nothing is skipped when it is asked for)
    est, n_used, err, tr, att = G.tomography(V[:5], delta, sched, norm_inf=norm_inf, stop=1, seed=55, vec_offset=9,
                                             trace=True)
    assert np.array_equal(est, fast[0][:5])
    for b in range(5):
        a = int(att[b])
        errs = np.linalg.norm(V[b] / np.linalg.norm(V[b]) - tr[b, :a], ord=np.inf if norm_inf else 2, axis=1)
        assert (errs[:-1] > delta).all() and errs[-1] <= delta


@pytest.mark.parametrize("d,delta,norm_inf", [(5000, 0.3, 0), (20000, 0.2, 0), (9000, 0.05, 1)])
def test_skipping_hopeless_attempts_changes_nothing_long_vectors(cuda, d, delta, norm_inf):
    """the same for vectors beyond the fused path (block / top kernels)"""
    from oracle import quantum_utility as qu
    from sqlearn import _cabi
    rs = np.random.RandomState(d)
    V = rs.uniform(-1, 1, (3, d))
    V[1] *= 4.0
    sched = qu.schedule(qu.tomography_N(d, delta, "inf" if norm_inf else "L2")).astype(float)
    try:
        _cabi.check(_cabi.lib.sq_tomography_tune(7, 0.0), "tune")
        plain = _big(V, delta, sched, norm_inf=norm_inf, stop=1, seed=66, vec_offset=3)
    finally:
        _cabi.check(_cabi.lib.sq_tomography_tune(7, 1.0), "tune")
    fast = _big(V, delta, sched, norm_inf=norm_inf, stop=1, seed=66, vec_offset=3)
    for a, b in zip(fast, plain):
        assert np.array_equal(a, b)
    assert (fast[2] <= delta).all() and (fast[1] < sched[-1]).all()
