"""Helpers shared by the CPU and GPU test modules (kept importable from the repo root on the GPU box)."""
import numpy as np


def pnp_case(rng, pn=9, noise=0.0):
    """A LINEMOD-like PnP problem: intrinsics, object keypoints, true [R|t], (noisy) projections, SPD covariances."""
    import cv2
    K = np.array([[572.4114, 0, 325.2611], [0, 573.57043, 242.04899], [0, 0, 1.0]])
    X = rng.uniform(-0.1, 0.1, (pn, 3))
    rvec = rng.normal(0, 0.9, 3)
    t = np.array([rng.uniform(-0.2, 0.2), rng.uniform(-0.15, 0.15), rng.uniform(0.5, 1.3)])
    R, _ = cv2.Rodrigues(rvec)
    P = X @ R.T + t
    x = np.stack([K[0, 0] * P[:, 0] / P[:, 2] + K[0, 2], K[1, 1] * P[:, 1] / P[:, 2] + K[1, 2]], 1)
    A = rng.normal(0, 1.5, (pn, 2, 2))
    cov = (A @ A.transpose(0, 2, 1) + 0.2 * np.eye(2)).astype(np.float32)
    return K, X, np.concatenate([R, t[:, None]], 1), x + rng.normal(0, noise, x.shape), cov
