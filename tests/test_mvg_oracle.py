"""The K8 checkers against each other (CPU): the reference's RANSAC scoring code compiled through
oracle/eigen_shim_mvg (its JacobiSVD restates Eigen's algorithm) vs the numpy / LAPACK restatement."""
import numpy as np
import pytest

from oracle import mvg_oracle as mo

needs_ref = pytest.mark.skipif(not mo.have_ref(), reason="oracle/_ref/libspectavi_ref_mvg.so not built")


@needs_ref
@pytest.mark.parametrize("n", [3, 4, 9])
def test_restated_jacobi_svd_is_an_svd(n):
    rng = np.random.RandomState(n)
    for trial in range(100):
        a = rng.randn(n, n) * 10.0 ** rng.randint(-3, 4)
        if trial % 4 == 0:
            a[:, -1] = a[:, 0] - a[:, 1]          # rank deficient (a fundamental matrix has rank 2)
        u, s, v = mo.ref_jacobi_svd(a)
        scale = np.abs(a).max()
        assert np.abs(u @ np.diag(s) @ v.T - a).max() < 1e-13 * scale
        assert np.abs(u.T @ u - np.eye(n)).max() < 1e-13 and np.abs(v.T @ v - np.eye(n)).max() < 1e-13
        assert (np.diff(s) <= 0).all() and (s >= 0).all()
        assert np.abs(s - np.linalg.svd(a, compute_uv=False)).max() < 1e-12 * scale


@needs_ref
def test_seven_point_reference_vs_numpy():
    # test/test_mvg.py:125-138: every solution satisfies x'^T F x = 0 on the seven matches; both solvers find the same set
    rng = np.random.RandomState(1)
    for _ in range(50):
        x, xp = rng.randn(7, 2), rng.randn(7, 2)
        Fr, Fn = mo.ref_seven_point(x, xp), mo.seven_point(x, xp)
        assert len(Fr) in (1, 2, 3) and len(Fr) == len(Fn)
        h = lambda a: np.hstack([a, np.ones((7, 1))])
        for F in Fr:
            assert np.abs(np.sum((h(xp) @ F) * h(x), axis=1)).max() < 1e-10 * np.abs(F).max()
            Fu = F / np.linalg.norm(F)
            assert min(min(np.abs(Fu - G / np.linalg.norm(G)).max(), np.abs(Fu + G / np.linalg.norm(G)).max()) for G in Fn) < 1e-6


@needs_ref
@pytest.mark.parametrize("seed,npt,outliers,noise,err_allowed", [(0, 200, .3, 0., 1e-6), (1, 500, .5, 1e-4, 3.35e-4), (2, 64, .2, 1e-3, 5e-3)])
def test_process_fundamental_matrix_reference_vs_lapack(seed, npt, outliers, noise, err_allowed):
    rng = np.random.RandomState(seed)
    x0, x1, E = mo.synthetic_two_view(rng, npt, outliers, noise)
    F = np.vstack([mo.hypotheses(rng, x0, x1, 30, use_ref=True), E.reshape(1, 9), rng.randn(3, 9)])
    ok, cnt, cams, masks, _ = mo.ref_process(x0, x1, F, required=.4, err_allowed=err_allowed, find_best=True, sv_ratio=3e-2)
    assert ok[-4] == 1 and cnt[-4] >= (1 - outliers) * npt * (0.9 if noise else 1.0) - 1     # the true E explains the inliers
    assert (ok[-3:] == 0).all()                                                          # random matrices fail the SVD gate
    for f in range(F.shape[0]):
        cand = mo.essential_cameras(F[f].reshape(3, 3), 3e-2)
        if cand is None:
            assert ok[f] == 0
            continue
        if not ok[f]:
            continue
        assert any(mo.same_camera(cams[f], P) for P in cand)
        m, decided = mo.inlier_mask(cams[f], x0, x1, err_allowed)
        assert (m[decided] == masks[f].astype(bool)[decided]).all()
        assert abs(int(m.sum()) - int(cnt[f])) <= int((~decided).sum())
        # the reference keeps the FIRST camera with the strictly largest count
        counts = [int(mo.inlier_mask(P, x0, x1, err_allowed)[0].sum()) for P in cand]
        assert cnt[f] >= max(counts) - int((~decided).sum()) - 1
