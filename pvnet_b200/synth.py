"""Synthetic mask + vertex fields for tests and bench (no dataset, no network).

The field generator restates what the reference's numpy `compute_vertex` does
(/root/reference/tools/demo.py:58-71, simulation.py:7-20): for every foreground
pixel and keypoint the unit vector pixel -> keypoint (norms below 1e-3 get
+1e-3).  On top of that SURVEY.md section 8(d) asks for per-pixel angular noise
and junk background vectors so that mask bugs show up.
"""
from __future__ import annotations

import os

import numpy as np

H, W = 480, 640
_GOLDEN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def cat_fixture():
    """configs[0]: the reference demo's cat mask (2289 fg px) and its 9 projected
    keypoints; stored in tests/golden/cat_fixture.npz by tests/golden/make_golden.py."""
    z = np.load(os.path.join(_GOLDEN, "cat_fixture.npz"))
    mask = np.unpackbits(z["mask_bits"])[: H * W].reshape(H, W).astype(np.int64)
    return mask, z["kp2d"].astype(np.float64)


def unit_field(mask, kps, sigma=0.0, rng=None, junk_background=True):
    """mask [h,w] (nonzero = fg), kps [vn,2] (x,y) -> vertex [h,w,vn,2] float32."""
    h, w = mask.shape
    vn = kps.shape[0]
    ys, xs = np.nonzero(mask)
    d = kps[None, :, :] - np.stack([xs, ys], 1)[:, None, :].astype(np.float64)  # [tn,vn,2]
    nrm = np.linalg.norm(d, axis=2, keepdims=True)
    nrm[nrm < 1e-3] += 1e-3
    d = d / nrm
    if sigma > 0:
        ang = rng.normal(0.0, sigma, size=d.shape[:2])
        c, s = np.cos(ang), np.sin(ang)
        d = np.stack([c * d[..., 0] - s * d[..., 1], s * d[..., 0] + c * d[..., 1]], -1)
    if junk_background:
        if rng is None:
            rng = np.random.default_rng(0)
        out = rng.standard_normal((h, w, vn, 2), dtype=np.float32)
    else:
        out = np.zeros((h, w, vn, 2), np.float32)
    out[ys, xs] = d.astype(np.float32)
    return out


def blob_mask(rng, frac=0.03, h=H, w=W):
    """One rotated ellipse covering ~frac of the image, with a ragged border."""
    area = frac * h * w
    ratio = rng.uniform(0.5, 1.0)
    a = np.sqrt(area / (np.pi * ratio))
    bb = a * ratio
    a, bb = min(a, 0.49 * w), min(bb, 0.49 * h)
    cx = rng.uniform(min(a, w / 2) , max(w - a, w / 2))
    cy = rng.uniform(min(a, h / 2), max(h - a, h / 2))
    th = rng.uniform(0, np.pi)
    yy, xx = np.mgrid[0:h, 0:w]
    u = (xx - cx) * np.cos(th) + (yy - cy) * np.sin(th)
    v = -(xx - cx) * np.sin(th) + (yy - cy) * np.cos(th)
    r = (u / a) ** 2 + (v / bb) ** 2
    m = r < 1.0 + 0.08 * rng.standard_normal((h, w))
    return m.astype(np.int64)


def keypoints_in_bbox(rng, mask, vn, margin=0.2):
    ys, xs = np.nonzero(mask)
    if len(xs) == 0:
        return rng.uniform([0, 0], [mask.shape[1], mask.shape[0]], size=(vn, 2))
    x0, x1, y0, y1 = xs.min(), xs.max(), ys.min(), ys.max()
    mx, my = margin * (x1 - x0 + 1), margin * (y1 - y0 + 1)
    return rng.uniform([x0 - mx, y0 - my], [x1 + mx, y1 + my], size=(vn, 2))


def make_batch(b, vn=9, frac=0.03, sigma=0.05, seed=1234, h=H, w=W):
    """configs[1]-style batch. Returns mask [b,h,w] int64, vertex [b,h,w,vn,2] f32, kps [b,vn,2]."""
    masks = np.zeros((b, h, w), np.int64)
    verts = np.zeros((b, h, w, vn, 2), np.float32)
    kps = np.zeros((b, vn, 2))
    for i in range(b):
        rng = np.random.default_rng(seed + i)
        masks[i] = blob_mask(rng, frac, h, w)
        kps[i] = keypoints_in_bbox(rng, masks[i], vn)
        verts[i] = unit_field(masks[i], kps[i], sigma, rng)
    return masks, verts, kps


def as_net_output(vertex, seg_dim=2, device=None, pin=False):
    """Lay vertex [b,h,w,vn,2] out the way Resnet18_8s emits it
    (/root/reference/lib/networks/model_repository_orig.py:76-80 and the
    permute/view at tools/train_linemod.py:118-120): channels 2: of one
    [b, seg_dim+2vn, h, w] tensor, returned as the NON-contiguous [b,h,w,vn,2] view."""
    import torch

    v = torch.as_tensor(vertex)
    b, h, w, vn, _ = v.shape
    full = torch.zeros((b, seg_dim + 2 * vn, h, w), dtype=torch.float32, device=device,
                       pin_memory=bool(pin and device is None))
    full[:, seg_dim:] = v.reshape(b, h, w, 2 * vn).permute(0, 3, 1, 2).to(full.device)
    view = full[:, seg_dim:].permute(0, 2, 3, 1).view(b, h, w, vn, 2)
    return view


def philox_note():
    """The index / selection streams are produced by the product's own
    counter-based generator (pvv_make_idxs / pvv_make_selection in include/pvvote.h)
    and *materialised* for the oracle -- see pvnet_b200.ransac_voting_gpu.make_idxs."""


def vanishing_point_field(rng, tn, vn, noise=0.0):
    """Pixels and unit directions pointing at homogeneous points V[vi] = (X, Y, Z); Z = 0 is a point at infinity."""
    coords = np.stack([rng.integers(200, 440, tn), rng.integers(100, 380, tn)], 1).astype(np.float32)
    V = np.stack([rng.uniform(-2000, 2000, vn), rng.uniform(-2000, 2000, vn), rng.choice([1.0, 2.5, 0.0], vn)], 1)
    V[0, 2], V[1, 2] = 0.0, 1.0
    diff = V[None, :, :2] - coords[:, None, :] * V[None, :, 2:3]
    ang = np.arctan2(diff[..., 1], diff[..., 0]) + rng.normal(0, noise, diff.shape[:2]) if noise else None
    direct = np.stack([np.cos(ang), np.sin(ang)], -1) if noise else diff / np.linalg.norm(diff, axis=-1, keepdims=True)
    return coords, direct.astype(np.float32), V


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
