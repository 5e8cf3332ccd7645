"""CPU-side checks of the product's host layer: the C-ABI library loads and
exports every symbol include/sqb200.h declares (no compute calls), and the
integer/host arithmetic of sqlearn.QuantumUtility equals the reference's
(golden fixtures)."""
import os
import re

import numpy as np
import pytest

from conftest import ROOT


def _declared_functions():
    src = open(os.path.join(ROOT, "include", "sqb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(sq_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(lib):
    names = _declared_functions()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib.lib, n), f"{n} declared in sqb200.h but missing from libsqb200.so"
        assert n in lib.SIGNATURES, f"{n} has no ctypes signature in sqlearn/_cabi.py"
    assert set(lib.SIGNATURES) == set(names)
    assert lib.lib.sq_abi_version() == 1


def test_argument_errors_do_not_need_a_gpu(lib):
    # validation happens before any CUDA call
    rc = lib.lib.sq_tomography_batched(None, 1, 4, 0.1, 0, 0, None, 1, 0, 0, None, None, None, None, None, None)
    assert rc == 1
    assert b"null pointer" in lib.lib.sq_last_error()
    with pytest.raises(lib.SqError):
        lib.check(rc, "sq_tomography_batched")


def test_host_schedule_and_counts_match_reference(golden_det):
    from sqlearn.QuantumUtility import Utility as U
    for d, delta, norm, N in golden_det["N"]:
        assert U.tomography_num_measurements(d, delta, norm) == N
    for N, fmi, sched in golden_det["schedule"]:
        assert [int(x) for x in U.tomography_schedule(N, fmi)] == sched
    for g, Q in golden_det["median_Q"]:
        assert U.median_evaluation_Q(g) == Q
    for e, M in golden_det["ae_M"]:
        assert U.amplitude_estimation_M(e) == M
    for e, g, m in golden_det["pe_m"]:
        assert U.phase_estimation_m(e, g) == m
    for v, j, out in golden_det["check_division"]:
        assert U.check_division(v, j) == out
    for x, w, back in golden_det["wrap"]:
        assert U.wrapper_phase_est_arguments(x) == w
        assert U.unwrap_phase_est_arguments(w / (0.01 + np.pi), 0.01) == back


def test_consistent_bucket_closed_form_equals_reference_grid(golden_det):
    from sqlearn.QuantumUtility import Utility as U
    for eps, gamma, pe, want in golden_det["cpe_bucket"]:
        n = U.phase_estimation_m(eps, gamma)
        C = gamma / n
        dp = eps * C / 2
        shift = int(np.floor(2 / C) / 2) + 1
        assert U._consistent_bucket(pe, eps, shift, dp) == want


def test_consistent_bucket_dense_sweep():
    """closed-form cell lookup == np.arange + bisect, on many outcomes"""
    from oracle import quantum_utility as oq
    from sqlearn.QuantumUtility import Utility as U
    rng = np.random.default_rng(0)
    for eps, gamma in [(0.05, 0.1), (0.013, 0.7), (0.2, 0.3)]:
        n = U.phase_estimation_m(eps, gamma)
        C = gamma / n
        dp = eps * C / 2
        shift = int(np.floor(2 / C) / 2) + 1
        for pe in np.concatenate([rng.random(300), np.arange(64) / 64]):
            assert U._consistent_bucket(pe, eps, shift, dp) == oq.consistent_bucket(pe, eps, gamma)[0]


def test_estimator_parameter_checks():
    from sqlearn.cluster import qMeans_
    est = qMeans_(n_clusters=3, delta=0.5, algorithm='full')
    assert est.get_params()["delta"] == 0.5
    est.set_params(n_clusters=4)
    assert est.n_clusters == 4
    X = np.zeros((2, 2))
    with pytest.raises(ValueError, match="n_samples=2 should be >= n_clusters=4"):
        est._check_params(X)
    with pytest.raises(ValueError, match="n_init should be > 0"):
        qMeans_(n_init=0)._check_params(np.zeros((10, 2)))
    with pytest.raises(ValueError, match="Algorithm must be"):
        qMeans_(algorithm="x")._check_params(np.zeros((10, 2)))
    e = qMeans_(n_clusters=2)
    e._check_params(np.zeros((10, 2)))
    assert e._algorithm == "elkan"  # the reference's 'auto' resolves to elkan (D3)


def test_kmeans_plusplus_consumes_random_state_like_reference():
    from oracle import refload
    if not refload.available():
        pytest.skip("reference mount absent")
    import warnings
    from sqlearn.cluster._dmeans import _kmeans_plusplus
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        dm = refload.load_dmeans()
    rs = np.random.RandomState(0)
    X = rs.randn(500, 6)
    xs = np.einsum("ij,ij->i", X, X)
    c_ref, i_ref = dm._kmeans_plusplus(X, 7, xs, np.random.RandomState(42))
    c_me, i_me = _kmeans_plusplus(X, 7, xs, np.random.RandomState(42))
    assert list(i_ref) == list(i_me)
    np.testing.assert_array_equal(c_ref, c_me)


def test_committed_bench_line_has_the_contract_keys():
    """the bench line committed under profiles/ carries every key the measurement contract names"""
    import json
    import os
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "r01_bench_final_1gpu.json")
    line = json.load(open(path))
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                "vs_baseline", "dtype", "data", "config", "clocks", "e2e", "gpu_launches", "roofline", "cpu_baseline"):
        assert key in line, key
    assert line["n_gpus"] == 1 and line["dtype"] == "f64" and line["vs_baseline"] is None
    assert set(("bound", "achieved", "peak", "unit", "frac", "traffic")) <= set(line["roofline"])
    assert abs(line["roofline"]["frac"] - line["roofline"]["achieved"] / line["roofline"]["peak"]) < 1e-9
    assert set(("value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step")) <= set(line["e2e"])
    assert set(("value", "unit", "cores", "kind", "sample")) <= set(line["cpu_baseline"])
    assert line["gpu_launches"] > 0 and "workload" in line["config"]
    assert not set(line["clocks"]["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}


def test_mle_dimension_matches_stock_scikit_learn():
    """n_components='mle' (reference: _infer_dimension, _qPCA.py:101-110, Minka's rule): the host restatement picks
    the rank the installed scikit-learn picks"""
    from sklearn.decomposition._pca import _infer_dimension as stock
    from sqlearn.decomposition._qPCA import _infer_dimension
    rs = np.random.RandomState(0)
    for n, d, r in [(200, 10, 3), (500, 30, 7), (100, 8, 2), (1000, 50, 10), (60, 6, 5)]:
        X = rs.randn(n, r) @ rs.randn(r, d) * 3 + rs.randn(n, d) * 0.3
        spectrum = np.linalg.svd(X - X.mean(axis=0), compute_uv=False) ** 2 / (n - 1)
        assert _infer_dimension(spectrum, n) == stock(spectrum, n)
