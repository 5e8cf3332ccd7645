"""The workspace-free E-step entry point (sq_qmeans_assign_noalloc).

It exists for callers that may not allocate device memory themselves - e.g. a
plugin living inside another framework's memory pool, or a serving loop that
cannot tolerate cudaMalloc's device-wide synchronisation - and keeps its
scratch in a fixed array inside the library.  Because that array is shared by
all calls and never cleared as a whole, the test makes two different calls
back to back and checks each against the workspace-taking entry point: stale
contents from the first call must not leak into the second.  Collected last
on purpose (file name), so a fail-fast run has already executed everything
else."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _call_noalloc(X, C, mode, delta, seed):
    import torch
    from sqlearn import _cabi, _device as dv
    L = _cabi.lib
    Xd = torch.as_tensor(np.ascontiguousarray(X)).cuda()
    Cd = torch.as_tensor(np.ascontiguousarray(C)).cuda()
    n, d = Xd.shape
    k = Cd.shape[0]
    xn = torch.empty(n, dtype=torch.float64, device="cuda")
    _cabi.check(L.sq_row_sqnorms(Xd.data_ptr(), n, d, xn.data_ptr(), dv.stream_ptr()), "norms")
    labels = torch.full((n,), -7, dtype=torch.int32, device="cuda")
    mind = torch.empty(n, dtype=torch.float64, device="cuda")
    inertia = torch.empty(1, dtype=torch.float64, device="cuda")
    nfix = torch.zeros(1, dtype=torch.int32, device="cuda")
    rc = L.sq_qmeans_assign_noalloc(Xd.data_ptr(), xn.data_ptr(), n, d, Cd.data_ptr(), k, mode, float(delta), seed, 0,
                                    labels.data_ptr(), mind.data_ptr(), inertia.data_ptr(), nfix.data_ptr(),
                                    dv.stream_ptr())
    torch.cuda.synchronize()
    return rc, labels.cpu().numpy(), mind.cpu().numpy(), float(inertia.item()), int(nfix.item())


def test_noalloc_matches_workspace_entry_point_twice_in_a_row(cuda):
    import gpu_util as G
    rs = np.random.RandomState(123)
    # call 1: exact mode with many exact ties (fills the near-tie list and its counter)
    X1 = np.zeros((5000, 6))
    C1 = np.vstack([np.eye(6), -np.eye(6)])
    # call 2: a different shape, few near ties; then a noisy (delta-window) call on the same data
    C2 = rs.randn(200, 33) * 2
    X2 = C2[rs.randint(0, 200, 3000)] + rs.randn(3000, 33)
    for X, C, mode, delta in [(X1, C1, 0, 0.0), (X2, C2, 0, 0.0), (X2, C2, 1, 6.0)]:
        rc, lab, mind, inertia, nfix = _call_noalloc(X, C, mode, delta, seed=77)
        assert rc == 0
        want_lab, want_mind, want_inertia, want_fix = G.assign(X, C, mode=mode, delta=delta, seed=77)
        assert np.array_equal(lab, want_lab)                      # same seed: identical, ties included
        np.testing.assert_allclose(mind, want_mind, rtol=1e-10, atol=1e-12)
        assert inertia == pytest.approx(want_inertia, rel=1e-10)
        assert nfix == want_fix


def test_noalloc_ipe_mode_and_wide_window(cuda):
    """mode 2 (the IPE tiles live in what the fixed part leaves of the 2 MiB: a smaller grid, same draws) and the
    delta-window mode with more than 256 centroids (row-minimum pre-pass), twice in a row each, plus the largest
    shape the entry point takes"""
    import gpu_util as G
    rs = np.random.RandomState(321)
    C3 = rs.randn(40, 16)
    X3 = C3[rs.randint(0, 40, 2500)] + 0.3 * rs.randn(2500, 16)
    C4 = rs.randn(300, 24) * 1.5
    X4 = C4[rs.randint(0, 300, 4000)] + rs.randn(4000, 24)
    for X, C, mode, delta, seed in [(X3, C3, 2, 0.5, 5), (X4, C4, 1, 5.0, 6), (X3, C3, 2, 0.5, 7), (X4, C4, 1, 5.0, 8),
                                    (X4, C4, 2, 1.0, 9)]:
        rc, lab, mind, inertia, nfix = _call_noalloc(X, C, mode, delta, seed=seed)
        assert rc == 0
        want_lab, want_mind, want_inertia, want_fix = G.assign(X, C, mode=mode, delta=delta, seed=seed)
        assert np.array_equal(lab, want_lab)                      # same seed, same draws, whatever the grid
        np.testing.assert_allclose(mind, want_mind, rtol=1e-10, atol=1e-12)
        assert inertia == pytest.approx(want_inertia, rel=1e-10)
    # 65536 x 1024: the fixed part of the scratch is 1.33 MB, seven IPE tiles fit beside it
    C5 = rs.randn(1024, 8)
    X5 = C5[rs.randint(0, 1024, 65536)] + 0.2 * rs.randn(65536, 8)
    for mode, delta in [(0, 0.0), (1, 0.3), (2, 0.3)]:
        rc, lab, mind, inertia, nfix = _call_noalloc(X5, C5, mode, delta, seed=11)
        assert rc == 0
        want_lab, want_mind, want_inertia, _ = G.assign(X5, C5, mode=mode, delta=delta, seed=11)
        assert np.array_equal(lab, want_lab)
        assert inertia == pytest.approx(want_inertia, rel=1e-10)


def test_noalloc_rejects_sizes_beyond_its_scratch(cuda, lib):
    rc, *_ = _call_noalloc(np.zeros((65537, 2)), np.ones((3, 2)), 0, 0.0, seed=1)
    assert rc == 4
    assert b"65536" in lib.lib.sq_last_error()
    rc, *_ = _call_noalloc(np.zeros((10, 2)), np.ones((1025, 2)), 0, 0.0, seed=1)
    assert rc == 4
    rc, lab, *_ = _call_noalloc(np.zeros((65536, 2)) + 0.1, np.array([[0.0, 0.0], [5.0, 5.0]]), 0, 0.0, seed=1)
    assert rc == 0 and (lab == 0).all()
