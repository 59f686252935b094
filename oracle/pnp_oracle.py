"""CPU restatement of the reference's uncertainty-driven PnP (TEST INFRASTRUCTURE ONLY -- see oracle/README.md).

  weights_from_covariance   <- Evaluator.evaluate_uncertainty, lib/utils/evaluation_utils.py:169-181
  uncertainty_pnp           <- lib/utils/extend_utils/extend_utils.py:63-114 (start pose: OpenCV P3P on the four
                               keypoints with the largest wxx + wxy) + the Ceres problem of
                               lib/utils/extend_utils/src/uncertainty_pnp.cpp:7-92

Third-party arithmetic that is NOT under /root/reference and could not be built here (parity unpinned for this row):
  * Ceres Solver 1.14 (prebuilt lib/utils/extend_utils/lib/libceres.so.1.14.0; its own dependencies -- glog, SuiteSparse,
    OpenBLAS, CXSparse -- are absent from this image): Levenberg-Marquardt trust region, DENSE_SCHUR, AutoDiff, default
    stopping rules (50 iterations, function tolerance 1e-6).  Restated here as scipy.optimize.least_squares(method="lm")
    (MINPACK's LM) on the same residuals, run to tight tolerances: both find the minimiser of the same cost; Ceres may stop
    up to its function tolerance short of it.
  * OpenCV 3.4.2 solvePnP(SOLVEPNP_P3P) (environment.yml): called here through the installed cv2 (4.x), as the reference
    calls it.  It only provides the start pose.
The reference holds no test or golden vector for this function; tests pin the restatement through known-answer cases
(exact projections of a known pose are recovered; the gradient of the cost vanishes at the returned pose).
"""
from __future__ import annotations

import numpy as np


def weights_from_covariance(covar):
    """evaluation_utils.py:169-181: [pn,2,2] covariances -> [pn,3] (wxx, wxy, wyy) of inv(sqrtm(cov))."""
    import scipy.linalg
    covar = np.asarray(covar)
    cov_invs = []
    for vi in range(covar.shape[0]):
        if covar[vi, 0, 0] < 1e-6 or np.sum(np.isnan(covar)[vi]) > 0:
            cov_invs.append(np.zeros([2, 2]).astype(np.float32))
            continue
        cov_invs.append(np.linalg.inv(scipy.linalg.sqrtm(covar[vi])))
    weights = np.asarray(cov_invs).reshape([-1, 4])
    return np.real(weights[:, (0, 1, 3)])


def _rodrigues_rotate(rt, X):
    """ceres::AngleAxisRotatePoint (ceres/rotation.h), as used at uncertainty_pnp.cpp:21."""
    w = rt[:3]
    th2 = float(w @ w)
    if th2 > np.finfo(np.float64).eps:
        th = np.sqrt(th2)
        k = w / th
        c, s = np.cos(th), np.sin(th)
        return X * c + np.cross(k, X) * s + np.outer(X @ k, k) * (1.0 - c)
    return X + np.cross(w, X)


def residuals(rt, pts2d, pts3d, w, K):
    """uncertainty_pnp.cpp:18-33: r_i = W_i (proj(R X_i + t) - x_i), stacked [2*pn]."""
    P = _rodrigues_rotate(rt, pts3d) + rt[3:6]
    px = K[0, 0] * P[:, 0] / P[:, 2] + K[0, 2]
    py = K[1, 1] * P[:, 1] / P[:, 2] + K[1, 2]
    dx, dy = px - pts2d[:, 0], py - pts2d[:, 1]
    return np.stack([w[:, 0] * dx + w[:, 1] * dy, w[:, 1] * dx + w[:, 2] * dy], 1).reshape(-1)


def uncertainty_pnp(points_2d, weights_2d, points_3d, camera_matrix, return_info=False):
    """extend_utils.py:63-114.  Returns Rt [3,4] float64."""
    import cv2
    from scipy.optimize import least_squares
    pn = points_2d.shape[0]
    assert points_3d.shape[0] == pn and pn >= 4
    dist_coeffs = np.zeros(shape=[8, 1], dtype=np.float64)
    points_3d = points_3d.astype(np.float64)
    points_2d = points_2d.astype(np.float64)
    weights_2d = weights_2d.astype(np.float64)
    camera_matrix = camera_matrix.astype(np.float64)
    idxs = np.argsort(weights_2d[:, 0] + weights_2d[:, 1])[-4:]
    _, R_exp, t = cv2.solvePnP(np.expand_dims(points_3d[idxs, :], 0), np.expand_dims(points_2d[idxs, :], 0),
                               camera_matrix, dist_coeffs, None, None, False, flags=cv2.SOLVEPNP_P3P)
    if pn == 4:
        R, _ = cv2.Rodrigues(R_exp)
        return np.concatenate([R, t], axis=-1)
    init_rt = np.concatenate([R_exp, t], 0).reshape(-1)
    sol = least_squares(residuals, init_rt, args=(points_2d, points_3d, weights_2d, camera_matrix), method="lm",
                        xtol=1e-15, ftol=1e-15, gtol=1e-15, max_nfev=2000)
    R, _ = cv2.Rodrigues(sol.x[:3])
    Rt = np.concatenate([R, sol.x[3:, None]], axis=-1)
    if return_info:
        return Rt, {"cost": float(2.0 * sol.cost), "nfev": int(sol.nfev), "init_rt": init_rt}
    return Rt


def evaluate_uncertainty_pose(mean_pts2d, covar, points_3d, K):
    """The pose part of Evaluator.evaluate_uncertainty (evaluation_utils.py:164-183)."""
    return uncertainty_pnp(mean_pts2d, weights_from_covariance(covar), points_3d, K)


# ---- the reference's own cost functor (oracle/_ref/ref_pnp_functor.so, built by oracle/build.py:build_ref_pnp) ----------
_REF_PNP = None


def ref_cost_available() -> bool:
    import os
    return os.path.exists(os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "ref_pnp_functor.so"))


def ref_cost(rt, pts2d, pts3d, w, K):
    """Residuals [2*pn] and Jacobian [2*pn, 6] of the REFERENCE's ReprojectionErrorArray (uncertainty_pnp.cpp:7-33) at the
    pose rt = (angle-axis, translation), differentiated by the reference's own ceres::Jet."""
    import ctypes
    import os
    global _REF_PNP
    if _REF_PNP is None:
        _REF_PNP = ctypes.CDLL(os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "ref_pnp_functor.so"))
        dp = ctypes.POINTER(ctypes.c_double)
        _REF_PNP.ref_pnp_cost.argtypes = [dp, dp, dp, dp, dp, ctypes.c_int, dp, dp]
    c = lambda a: np.ascontiguousarray(a, np.float64)
    pts2d, pts3d, w, K, rt = c(pts2d), c(pts3d), c(w), c(K), c(rt)
    pn = pts2d.shape[0]
    res, jac = np.zeros(2 * pn), np.zeros((2 * pn, 6))
    p = lambda a: a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))
    _REF_PNP.ref_pnp_cost(p(pts2d), p(pts3d), p(w), p(K), p(rt), pn, p(res), p(jac))
    return res, jac
