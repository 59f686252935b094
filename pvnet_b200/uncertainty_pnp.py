"""Uncertainty-driven PnP on the GPU -- the consumer of estimate_voting_distribution_with_mean.

Mirrors
    uncertainty_pnp(points_2d, weights_2d, points_3d, camera_matrix)   lib/utils/extend_utils/extend_utils.py:63-114
    the pose part of Evaluator.evaluate_uncertainty                     lib/utils/evaluation_utils.py:164-183
over the C ABI entry pvv_uncertainty_pnp (include/pvvote.h, csrc/pvv_pnp.cu): one launch solves the whole batch, the
keypoints and covariances stay on the device (the reference copies them to the host and solves image by image through
cffi + Ceres).  No CPU fallback: a missing library or device raises.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib


def _dev_f64(x, device):
    return torch.as_tensor(np.asarray(x, np.float64) if not isinstance(x, torch.Tensor) else x).to(device=device, dtype=torch.float64).contiguous()


CERES_DEFAULT_STOP = (50, 1e-6)   # max_num_iterations, function_tolerance: what the reference's Solver::Options() holds


def uncertainty_pnp_batch(points_2d, cov_or_weights, points_3d, camera_matrix, weights=False, return_info=False, stop=None):
    """points_2d [b,vn,2], cov_or_weights [b,vn,2,2] covariances (weights=False: evaluate_uncertainty's conversion is done
    on the device) or [b,vn,3] (wxx,wxy,wyy; weights=True), float32 or float64 CUDA tensors; points_3d [vn,3] or [b,vn,3];
    camera_matrix [3,3] or [b,3,3].  Returns poses [b,3,4] float64 on the same device (+ info [b,3]: cost, iterations,
    status).  `stop` = (max_iter, function_tolerance), e.g. CERES_DEFAULT_STOP, ends the iteration the way the reference's
    solver options do (uncertainty_pnp.cpp:85-90); None runs to convergence."""
    if not torch.cuda.is_available():
        raise RuntimeError("pvnet_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    if not (isinstance(points_2d, torch.Tensor) and points_2d.is_cuda):
        raise RuntimeError("points_2d must be a CUDA tensor")
    dev = points_2d.device
    b, vn, _ = points_2d.shape
    f64 = points_2d.dtype == torch.float64 or cov_or_weights.dtype == torch.float64
    dt = torch.float64 if f64 else torch.float32
    p2 = points_2d.to(dtype=dt).contiguous()
    cw = cov_or_weights.to(device=dev, dtype=dt).contiguous()
    if tuple(cw.shape) != ((b, vn, 3) if weights else (b, vn, 2, 2)):
        raise ValueError(f"expected {'weights [b,vn,3]' if weights else 'covariances [b,vn,2,2]'}, got {tuple(cw.shape)}")
    p3 = _dev_f64(points_3d, dev)
    K = _dev_f64(camera_matrix, dev)
    if p3.shape[-2:] != (vn, 3) or K.shape[-2:] != (3, 3):
        raise ValueError("points_3d must be [vn,3] or [b,vn,3], camera_matrix [3,3] or [b,3,3]")
    p3_stride = vn * 3 if p3.dim() == 3 else 0
    k_stride = 9 if K.dim() == 3 else 0
    pose = torch.empty((b, 3, 4), dtype=torch.float64, device=dev)
    info = torch.empty((b, 3), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        mi, ft = (int(stop[0]), float(stop[1])) if stop is not None else (0, 0.0)
        st = _lib.lib().pvv_uncertainty_pnp_stop(p2.data_ptr(), cw.data_ptr(), 1 if weights else 0, 1 if f64 else 0,
                                                 p3.data_ptr(), p3_stride, K.data_ptr(), k_stride, b, vn, mi, ft,
                                                 pose.data_ptr(), info.data_ptr(),
                                                 C.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
    _lib.check(st, "pvv_uncertainty_pnp")
    return (pose, info) if return_info else pose


def uncertainty_pnp(points_2d, weights_2d, points_3d, camera_matrix):
    """The reference's per-image numpy signature (extend_utils.py:63-70): [pn,2], [pn,3] wxx,wxy,wyy, [pn,3], [3,3] ->
    Rt [3,4] float64 numpy."""
    pn = points_2d.shape[0]
    assert points_3d.shape[0] == pn and pn >= 4
    dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else None
    if dev is None:
        raise RuntimeError("pvnet_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    p2 = torch.as_tensor(np.asarray(points_2d, np.float64)[None]).to(dev)
    w = torch.as_tensor(np.asarray(weights_2d, np.float64)[None]).to(dev)
    return uncertainty_pnp_batch(p2, w, points_3d, camera_matrix, weights=True)[0].cpu().numpy()


def evaluate_uncertainty_pose(mean_pts2d, covar, points_3d, camera_matrix):
    """Pose part of Evaluator.evaluate_uncertainty (evaluation_utils.py:164-183) for a batch: mean_pts2d [b,vn,2] and
    covar [b,vn,2,2] straight from ransac_voting_layer_v3 / estimate_voting_distribution_with_mean."""
    return uncertainty_pnp_batch(mean_pts2d, covar, points_3d, camera_matrix, weights=False)
