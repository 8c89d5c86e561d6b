"""CPU checkers for K8, RANSAC hypothesis scoring (TEST INFRASTRUCTURE ONLY).

Two independent checkers:

* ``_ref/libspectavi_ref_mvg.so`` -- the reference's OWN code (src/RansacFitter.h:42-98
  process_fundamental_matrix, src/DltTriangulator.h, src/Camera.h:32-59, and the seven-point solver
  of src/FundamentalMatrixFitter.h) compiled where it lies through oracle/eigen_shim_mvg, whose
  JacobiSVD restates Eigen's published two-sided Jacobi algorithm (oracle/ref_mvg_driver.cpp).  The CUDA
  kernel performs the same operations in the same order, so results are compared with ``==``.
  Parity with a build against a real Eigen is UNPINNED (Eigen is absent from this image and its version
  is not pinned by the reference).
* numpy / LAPACK -- a restatement of the same functions on ``np.linalg.svd``, independent of the Jacobi
  restatement: inlier decisions must agree for every match whose reprojection error / depths are not
  within a stated relative margin of the decision boundary.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg import this module.
"""
import ctypes as ct
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_REF_SO = os.path.join(_HERE, "_ref", "libspectavi_ref_mvg.so")
_dp = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
_ip = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_ref = None


def have_ref():
    return os.path.exists(_REF_SO)


def ref():
    global _ref
    if _ref is None:
        if not have_ref():
            raise RuntimeError("oracle/_ref/libspectavi_ref_mvg.so not built (needs /root/reference at build time)")
        R = ct.CDLL(_REF_SO)
        R.ref_mvg_jacobi_svd.restype = None
        R.ref_mvg_jacobi_svd.argtypes = [_dp, ct.c_int, _dp, _dp, _dp]
        R.ref_mvg_seven_point.restype = ct.c_int
        R.ref_mvg_seven_point.argtypes = [_dp, _dp, _dp]
        R.ref_mvg_dlt_reprojection_error.restype = None
        R.ref_mvg_dlt_reprojection_error.argtypes = [_dp, _dp, ct.c_int, _dp, _dp, _dp, ct.c_void_p, ct.c_void_p]
        R.ref_mvg_process.restype = ct.c_double
        R.ref_mvg_process.argtypes = [_dp, _dp, ct.c_int, ct.c_double, ct.c_double, ct.c_int, ct.c_double, _dp, ct.c_int,
                                      ct.c_int, _ip, _ip, _dp, ct.c_void_p]
        _ref = R
    return _ref


# ------------------------------------------------------------------ the reference's own code
def ref_jacobi_svd(a):
    a = np.ascontiguousarray(a, np.float64)
    n = a.shape[0]
    sv, u, v = np.zeros(n), np.zeros((n, n)), np.zeros((n, n))
    ref().ref_mvg_jacobi_svd(a, n, sv, u, v)
    return u, sv, v


def ref_seven_point(x, xp):
    """src/Spectavi.cpp:14-36 on 7 x 2 inhomogeneous points: (nroots, 3, 3)."""
    dst = np.zeros((3, 9))
    n = ref().ref_mvg_seven_point(np.ascontiguousarray(x, np.float64), np.ascontiguousarray(xp, np.float64), dst)
    return dst[:n].reshape(n, 3, 3).copy()


def ref_process(x0, x1, F, required=.9, err_allowed=.5, find_best=True, sv_ratio=3e-2, nthreads=1, want_masks=True):
    """process_fundamental_matrix for every F: (ok int32 [nF], ninlier int32 [nF], cams [nF, 12], masks uint8
    [nF, npt] or None, seconds)."""
    x0 = np.ascontiguousarray(x0, np.float64)
    x1 = np.ascontiguousarray(x1, np.float64)
    F = np.ascontiguousarray(F, np.float64).reshape(-1, 9)
    nF, npt = F.shape[0], x0.shape[0]
    ok, cnt, cams = np.zeros(nF, np.int32), np.zeros(nF, np.int32), np.zeros((nF, 12))
    masks = np.zeros((nF, npt), np.uint8) if want_masks else None
    secs = ref().ref_mvg_process(x0, x1, npt, required, err_allowed, int(find_best), sv_ratio, F, nF, nthreads, ok, cnt, cams,
                                 masks.ctypes.data if want_masks else None)
    return ok, cnt, cams, masks, secs


# ------------------------------------------------------------------ numpy / LAPACK restatement
def seven_point(x, xp):
    """FundamentalMatrixFitter (src/FundamentalMatrixFitter.h:108-240) on numpy: the two null vectors of the 7 x 9
    system, the cubic det(z F0 + (1 - z) F1) = 0, one F per real root.  (The cubic's coefficients are recovered
    from four evaluations instead of the reference's expanded polynomial: same polynomial.)"""
    x, xp = np.asarray(x, np.float64), np.asarray(xp, np.float64)
    A = np.stack([xp[:, 0] * x[:, 0], xp[:, 0] * x[:, 1], xp[:, 0], xp[:, 1] * x[:, 0], xp[:, 1] * x[:, 1], xp[:, 1],
                  x[:, 0], x[:, 1], np.ones(7)], axis=1)
    _, _, vt = np.linalg.svd(A)
    F0, F1 = vt[7].reshape(3, 3), vt[8].reshape(3, 3)
    zs = np.array([-1., 0., 1., 2.])
    dets = np.array([np.linalg.det(z * F0 + (1 - z) * F1) for z in zs])
    coef = np.linalg.solve(np.vander(zs, 4), dets)
    if abs(coef[0]) < 1e-14:
        return np.zeros((0, 3, 3))
    roots = np.roots(coef)
    out = [r.real * F0 + (1 - r.real) * F1 for r in roots if abs(r.imag) < 1e-9]
    return np.array(out).reshape(-1, 3, 3)


def essential_cameras(F, sv_ratio):
    """RansacFitter.h:45-56 + Camera.h:32-59 on LAPACK: None when the gate rejects, else the four 3 x 4 cameras.
    (Singular vectors of a repeated / zero singular value are defined up to rotation / sign, so the cameras agree
    with another SVD's only as a projective SET.)"""
    u, s, vt = np.linalg.svd(F)
    if abs(s[0] - s[1]) / (abs(s[0] + s[1]) / 2.) > sv_ratio:
        return None
    E = u @ np.diag([1., 1., 0.]) @ vt
    u, _, vt = np.linalg.svd(E)
    D = np.array([[0., 1., 0.], [-1., 0., 0.], [0., 0., 1.]])
    t = u[:, 2:3]
    Ra, Rb = u @ D @ vt, u @ D.T @ vt
    return [np.hstack([Ra, t]), np.hstack([Ra, -t]), np.hstack([Rb, t]), np.hstack([Rb, -t])]


def dlt_score(P1, x0, x1):
    """DltTriangulator (src/DltTriangulator.h:28-91) with P0 = [I | 0] on LAPACK, all matches at once:
    (reprojection error, distance to camera 0, distance to camera 1)."""
    P0 = np.hstack([np.eye(3), np.zeros((3, 1))])
    P1 = np.asarray(P1, np.float64).reshape(3, 4)
    a, b = x0[:, :2] / x0[:, 2:3], x1[:, :2] / x1[:, 2:3]
    A = np.stack([a[:, 0:1] * P0[2] - P0[0], a[:, 1:2] * P0[2] - P0[1], b[:, 0:1] * P1[2] - P1[0], b[:, 1:2] * P1[2] - P1[1]], axis=1)
    X = np.linalg.svd(A)[2][:, 3, :]
    r0, r1 = X @ P0.T, X @ P1.T
    err = np.linalg.norm(r0[:, :2] / r0[:, 2:3] - a, axis=1) + np.linalg.norm(r1[:, :2] / r1[:, 2:3] - b, axis=1)
    s1 = -1. if np.linalg.det(P1[:, :3]) < 0 else 1.
    d0 = r0[:, 2] / X[:, 3]
    d1 = s1 / np.sum(P1[:, 2] ** 2) * r1[:, 2] / X[:, 3]
    return err, d0, d1


def inlier_mask(P1, x0, x1, err_allowed, margin=1e-9):
    """(mask, decided): the inlier decision of RansacFitter.h:64-66 for camera P1 and which matches are decided
    beyond `margin` (relative distance of the error from the threshold, of the depths from zero)."""
    err, d0, d1 = dlt_score(P1, x0, x1)
    mask = (err <= err_allowed) & (d0 > 0) & (d1 > 0)
    decided = (np.abs(err - err_allowed) > margin * max(err_allowed, 1e-300)) & (np.abs(d0) > margin) & (np.abs(d1) > margin)
    return mask, decided


def same_camera(Pa, Pb, tol=1e-7):
    """3 x 4 cameras equal up to an overall sign (the projective freedom an SVD's sign choices leave)."""
    Pa, Pb = np.asarray(Pa).reshape(3, 4), np.asarray(Pb).reshape(3, 4)
    return min(np.abs(Pa - Pb).max(), np.abs(Pa + Pb).max()) < tol


def synthetic_two_view(rng, npt, outlier_frac=0.3, noise=0.0):
    """The scene of test/test_mvg.py:38-66 (two cameras looking at a randn point cloud) expressed with camera 0 at
    [I | 0]; returns homogeneous normalised matches (x0, x1) and the true essential matrix.  A fraction of the
    matches is replaced by random points (outliers)."""
    def rot(a, b):
        v = np.cross(a, b)
        sk = np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0]])
        return np.eye(3) + sk + sk @ sk / (1 + a @ b)
    ang = rng.randn(3) * 0.2
    R = rot(np.array([0., 0., 1.]), (lambda w: w / np.linalg.norm(w))(np.array([ang[0], ang[1], 1.])))
    t = rng.randn(3)
    t /= np.linalg.norm(t)
    X = np.hstack([rng.randn(npt, 2) * 2, 6 + rng.rand(npt, 1) * 4])
    x0 = X.copy()
    x1 = X @ R.T + t
    x0 = x0 / x0[:, 2:3] + np.hstack([rng.randn(npt, 2) * noise, np.zeros((npt, 1))])
    x1 = x1 / x1[:, 2:3] + np.hstack([rng.randn(npt, 2) * noise, np.zeros((npt, 1))])
    nout = int(outlier_frac * npt)
    if nout:
        sel = rng.choice(npt, nout, replace=False)
        x1[sel, :2] = rng.randn(nout, 2) * 0.3
    # arbitrary homogeneous scales, as the reference's callers pass (test_mvg.py:62-63)
    x0 = x0 * (0.5 + rng.rand(npt, 1))
    x1 = x1 * (0.5 + rng.rand(npt, 1))
    tx = np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]])
    return np.ascontiguousarray(x0), np.ascontiguousarray(x1), tx @ R


def hypotheses(rng, x0, x1, ntries, use_ref=False):
    """Fundamental-matrix hypotheses as fit_essential draws them (RansacFitter.h:193-207): 7 random matches per
    try, every real root of the seven-point cubic.  (nF, 9)."""
    a, b = x0[:, :2] / x0[:, 2:3], x1[:, :2] / x1[:, 2:3]
    out = []
    for _ in range(ntries):
        sel = rng.choice(x0.shape[0], 7, replace=False)
        Fs = ref_seven_point(a[sel], b[sel]) if use_ref else seven_point(a[sel], b[sel])
        out.extend(F.reshape(9) for F in Fs)
    return np.ascontiguousarray(np.array(out).reshape(-1, 9))
