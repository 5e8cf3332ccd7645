"""Edge cases of the public surface: odd inputs, error behaviour, degenerate shapes."""
import warnings

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_qmeans_input_validation_and_small_shapes(cuda):
    from sqlearn.cluster import qMeans_
    rs = np.random.RandomState(0)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        with pytest.raises(ValueError, match="Expected 2D array"):
            qMeans_(n_clusters=2, algorithm='full', delta=0).fit(np.zeros(5))
        with pytest.raises(ValueError, match="0 sample"):
            qMeans_(n_clusters=1, algorithm='full', delta=0).fit(np.zeros((0, 3)))
        with pytest.raises(ValueError, match="NaN"):
            X = rs.randn(20, 3)
            X[3, 1] = np.nan
            qMeans_(n_clusters=2, algorithm='full', delta=0).fit(X)
        with pytest.raises(ValueError, match="n_samples=3 should be >= n_clusters=5"):
            qMeans_(n_clusters=5, algorithm='full', delta=0).fit(rs.randn(3, 2))
        with pytest.raises(ValueError, match="does not match the number of clusters"):
            qMeans_(n_clusters=2, algorithm='full', delta=0, init=np.zeros((3, 2))).fit(rs.randn(10, 2))
        with pytest.raises(ValueError, match="does not match the number of features"):
            qMeans_(n_clusters=2, algorithm='full', delta=0, init=np.zeros((2, 5))).fit(rs.randn(10, 2))
        # lists, float32 and Fortran order are accepted (promoted to C-ordered fp64)
        X = rs.randn(50, 3)
        a = qMeans_(n_clusters=2, algorithm='full', delta=0, init=X[:2], n_init=1).fit(X)
        b = qMeans_(n_clusters=2, algorithm='full', delta=0, init=X[:2], n_init=1).fit(np.asfortranarray(X))
        c = qMeans_(n_clusters=2, algorithm='full', delta=0, init=X[:2], n_init=1).fit(X.tolist())
        assert a.labels_ == b.labels_ == c.labels_
        f = qMeans_(n_clusters=2, algorithm='full', delta=0, init=X[:2].astype(np.float32), n_init=1)
        f = f.fit(X.astype(np.float32))
        assert len(f.labels_) == 50 and f.cluster_centers_.dtype == np.float64
        # one sample / one cluster / one feature
        one = qMeans_(n_clusters=1, algorithm='full', delta=0, n_init=1).fit(np.array([[3.0]]))
        assert one.labels_ == [0] and one.cluster_centers_.tolist() == [[3.0]] and one.inertia_ == 0
        # copy_x=True leaves the caller's array alone
        X0 = X.copy()
        qMeans_(n_clusters=2, algorithm='full', delta=0, n_init=2, random_state=0).fit(X)
        assert np.array_equal(X, X0)
        # not fitted / wrong width at predict time
        with pytest.raises(Exception, match="not fitted"):
            qMeans_(n_clusters=2).predict(X)
        with pytest.raises(ValueError, match="expecting 3 features"):
            a.predict(rs.randn(4, 2))
        # predict with noise returns a list of int, transform the n x k distances
        p = a.predict(X, delta=0.5)
        assert isinstance(p, list) and isinstance(p[0], int) and len(p) == 50
        D = a.transform(X)
        assert D.shape == (50, 2) and np.array_equal(np.argmin(D, axis=1), np.asarray(a.labels_))
        assert a.fit_predict(X) == a.labels_
        # identical points and identical centres: exact ties are broken at random (reference: _dmeans.py:772)
        t = qMeans_(n_clusters=3, algorithm='full', delta=0, n_init=1, init=np.zeros((3, 2)), max_iter=1)
        t = t.fit(np.ones((3000, 2)))
        assert min(np.bincount(t.labels_, minlength=3)) > 800


def test_empty_cluster_keeps_previous_centre(cuda):
    from sqlearn.cluster import qMeans_
    X = np.vstack([np.zeros((20, 2)), np.ones((20, 2)) * 10])
    init = np.array([[0.0, 0.0], [10.0, 10.0], [100.0, 100.0]])
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        est = qMeans_(n_clusters=3, algorithm='full', delta=0, n_init=1, init=init, max_iter=3).fit(X)
    assert any("no sample" in str(x.message) for x in w)
    assert any("distinct clusters (2)" in str(x.message) for x in w)      # the reference's ConvergenceWarning (:1313-1319)
    np.testing.assert_allclose(est.cluster_centers_[2], [100.0, 100.0])
    assert set(est.labels_) == {0, 1}


def test_tomography_argument_errors(cuda, lib):
    import torch
    from sqlearn import QuantumUtility as Q
    from sqlearn import _device as dv
    v = np.ones(8) / np.sqrt(8)
    with pytest.raises(ValueError, match="norm must be"):
        Q.tomography(v, 0.1, norm='L1')
    with pytest.raises(ValueError, match="norm must be"):
        Q.real_tomography(v, delta=0.1, norm='bad')
    t = torch.ones((2, 8), dtype=torch.float64, device="cuda")
    sched = np.array([10.0])
    out = torch.empty_like(t)
    s1 = torch.empty(2, dtype=torch.float64, device="cuda")
    rc = lib.lib.sq_tomography_batched(t.data_ptr(), 2, 8, 0.0, 0, 0, sched.ctypes.data, 1, 0, 0, out.data_ptr(),
                                       s1.data_ptr(), s1.data_ptr(), None, None, dv.stream_ptr())
    assert rc == 1 and b"delta" in lib.lib.sq_last_error()
    rc = lib.lib.sq_tomography_batched(t.data_ptr(), 2, 8, 0.1, 0, 0, sched.ctypes.data, 500, 0, 0, out.data_ptr(),
                                       s1.data_ptr(), s1.data_ptr(), None, None, dv.stream_ptr())
    assert rc == 1 and b"sched_len" in lib.lib.sq_last_error()
    rc = lib.lib.sq_tomography_batched(t.data_ptr(), 2, 5000, 0.1, 0, 0, sched.ctypes.data, 1, 0, 0, out.data_ptr(),
                                       s1.data_ptr(), s1.data_ptr(), None, None, dv.stream_ptr())
    assert rc == 4 and b"fused-path limit" in lib.lib.sq_last_error()
    # d = 1: N = 36*1*ln(1)/delta^2 = 0 and the reference's np.geomspace(1, 0) raises; so does the mirror.
    with pytest.raises(ValueError, match="Geometric sequence"):
        Q.tomography(np.array([[-2.0], [0.5]]), 0.3)
    # with an explicit N the only possible estimate is +-1
    e = Q.tomography(np.array([[-2.0], [0.5]]), 0.3, N=50, incremental_measure=False)
    assert e.tolist() == [[-1.0], [1.0]]
    # B = 0
    assert Q.tomography(np.zeros((0, 4)), 0.3).shape == (0, 4)


def test_workspace_too_small_is_reported(cuda, lib):
    import torch
    from sqlearn import _device as dv
    X = torch.zeros((100, 4), dtype=torch.float64, device="cuda")
    C = torch.zeros((3, 4), dtype=torch.float64, device="cuda")
    xn = torch.zeros(100, dtype=torch.float64, device="cuda")
    lab = torch.zeros(100, dtype=torch.int32, device="cuda")
    inertia = torch.zeros(1, dtype=torch.float64, device="cuda")
    ws = torch.zeros(64, dtype=torch.uint8, device="cuda")
    rc = lib.lib.sq_qmeans_assign(X.data_ptr(), xn.data_ptr(), 100, 4, C.data_ptr(), 3, 0, 0.0, 0, 0, lab.data_ptr(),
                                  None, inertia.data_ptr(), None, ws.data_ptr(), 64, dv.stream_ptr())
    assert rc == 3 and b"workspace needs" in lib.lib.sq_last_error()
    rc = lib.lib.sq_qmeans_assign(X.data_ptr(), xn.data_ptr(), 100, 4, C.data_ptr(), 3, 1, 0.0, 0, 0, lab.data_ptr(),
                                  None, inertia.data_ptr(), None, ws.data_ptr(), 64, dv.stream_ptr())
    assert rc == 1 and b"delta > 0" in lib.lib.sq_last_error()


def test_chunked_upload_pageable_and_pinned(cuda):
    """stream_to_device: many chunks, both staging buffers reused, callbacks cover every row once and see the
    chunk's data; pinned input takes the direct path"""
    import torch
    from sqlearn import _device as dv
    rs = np.random.RandomState(0)
    X = rs.randn(1003, 17)
    for src in (X, torch.from_numpy(X.copy()).pin_memory().numpy()):
        seen, sums = [], []

        def on_chunk(out, r0, r1):
            seen.append((r0, r1))
            sums.append(out[r0:r1].sum())          # stream-ordered after the chunk's copy
        out = dv.stream_to_device(src, on_chunk, chunk_bytes=17 * 8 * 100)
        torch.cuda.synchronize()
        assert np.array_equal(out.cpu().numpy(), X)
        assert seen == [(r, min(r + 100, 1003)) for r in range(0, 1003, 100)]
        np.testing.assert_allclose([float(v) for v in sums], [X[a:b].sum() for a, b in seen], rtol=1e-12, atol=1e-12)
    assert dv.stream_to_device(np.zeros((0, 5))).shape == (0, 5)


@pytest.mark.parametrize("init", ["random", "k-means++", "array"])
@pytest.mark.parametrize("copy_x", [True, False])
def test_fit_leaves_the_callers_array_alone(cuda, init, copy_x):
    """copy_x=True: X untouched; copy_x=False: the reference centres in place and adds the mean back, so X comes
    back equal up to rounding (reference: _dmeans.py:1262-1269, :1311-1312)"""
    import warnings
    from sqlearn.cluster import qMeans_
    rs = np.random.RandomState(3)
    X = rs.randn(500, 6) + 5.0
    X0 = X.copy()
    ini = X[:4].copy() if init == "array" else init
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        qMeans_(n_clusters=4, init=ini, n_init=1, max_iter=3, delta=0.3, algorithm='full', copy_x=copy_x,
                true_distance_estimate=False, random_state=0).fit(X)
    if copy_x or init != "random":
        assert np.array_equal(X, X0)
    else:
        np.testing.assert_allclose(X, X0, rtol=1e-12, atol=1e-12)
