"""CPU oracle for the RANSAC voting path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this module; nothing under pvnet_b200/ does.

Control flow restates (paths relative to /root/reference):
  ransac_voting_layer_v3                  lib/ransac_voting_gpu_layer/ransac_voting_gpu.py:520-604
  estimate_voting_distribution_with_mean  lib/ransac_voting_gpu_layer/ransac_voting_gpu.py:333-406
  b_inv                                   lib/ransac_voting_gpu_layer/ransac_voting_gpu.py:503-518
over the kernels in oracle/pvv_oracle.c (ransac_voting_kernel.cu:11-49, 88-126).

Random draws are *injected*: the reference calls Tensor.random_/uniform_
(ransac_voting_gpu.py:544,553,352,367); here the caller passes `idxs`
(int32 [b,hn,vn,2] for v3, [b,rounds,hn,vn,2] for the distribution) and
`selection` (float32 [b,h,w] uniforms) explicitly, so that the product, this
oracle and the reference wrapper can be run on identical streams.

Deliberate, documented deviations (DESIGN.md "Reference behaviours"):
  * v3's while-loop re-votes the same idxs each round; the result
    is a function of round 0 only, so one round is evaluated.
  * the refinement sums and the covariance are accumulated in float64 from the
    same float32 terms; the reference's float32 sum order is library-defined.
    Parity on those outputs is 1e-4 relative (BASELINE.json north_star).
  * b_inv's whole-batch identity fallback for a singular ATA (:514-517) is
    applied per keypoint.
Parity pin: tests/golden/*.npz were produced by running the reference's own
Python wrapper over these kernels (tests/golden/make_golden.py); on the GPU box
the kernels are A/B-ed bitwise against the reference's CUDA build (oracle/_ref).
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

_f32p = ctypes.POINTER(ctypes.c_float)
_i32p = ctypes.POINTER(ctypes.c_int32)
_i64p = ctypes.POINTER(ctypes.c_int64)
_u8p = ctypes.POINTER(ctypes.c_uint8)


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liborc.so")
        if not os.path.exists(path):
            from . import build as _b  # type: ignore
            _b.build_oracle()
        L = ctypes.CDLL(path)
        L.orc_generate_hypothesis.argtypes = [_f32p, _f32p, _i32p, ctypes.c_int, ctypes.c_int, ctypes.c_int, _f32p]
        L.orc_voting_for_hypothesis.argtypes = [_f32p, _f32p, _f32p, _u8p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_float]
        L.orc_vote_counts.argtypes = [_f32p, _f32p, _f32p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_float, _i64p]
        L.orc_inlier_test.argtypes = [ctypes.c_float] * 7
        L.orc_generate_hypothesis_vanishing_point.argtypes = [_f32p, _f32p, _i32p, ctypes.c_int, ctypes.c_int, ctypes.c_int, _f32p]
        L.orc_voting_for_hypothesis_vanishing_point.argtypes = [_f32p, _f32p, _f32p, _u8p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_float]
        L.orc_inlier_test.restype = ctypes.c_int
        L.orc_num_threads.restype = ctypes.c_int
        _LIB = L
    return _LIB


def _p(a, t):
    return a.ctypes.data_as(t)


def _c(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


# ----------------------------------------------------------------------------
# kernels (reference layouts: direct [tn,vn,2], coords [tn,2], idxs/hyp [hn,vn,2])
# ----------------------------------------------------------------------------
def generate_hypothesis(direct, coords, idxs):
    direct, coords, idxs = _c(direct, np.float32), _c(coords, np.float32), _c(idxs, np.int32)
    tn, vn, _ = direct.shape
    hn = idxs.shape[0]
    hyp = np.empty((hn, vn, 2), np.float32)
    lib().orc_generate_hypothesis(_p(direct, _f32p), _p(coords, _f32p), _p(idxs, _i32p), tn, vn, hn, _p(hyp, _f32p))
    return hyp


def generate_hypothesis_vanishing_point(direct, coords, idxs):
    """ransac_voting_kernel.cu:170-227: homogeneous hypotheses [hn,vn,3]."""
    direct, coords, idxs = _c(direct, np.float32), _c(coords, np.float32), _c(idxs, np.int32)
    tn, vn, _ = direct.shape
    hn = idxs.shape[0]
    hyp = np.empty((hn, vn, 3), np.float32)
    lib().orc_generate_hypothesis_vanishing_point(_p(direct, _f32p), _p(coords, _f32p), _p(idxs, _i32p), tn, vn, hn,
                                                  _p(hyp, _f32p))
    return hyp


def voting_for_hypothesis_vanishing_point(direct, coords, hyp, inliers, thresh):
    """ransac_voting_kernel.cu:258-301, in place on `inliers` (uint8 [hn,vn,tn], pre-zeroed)."""
    direct, coords, hyp = _c(direct, np.float32), _c(coords, np.float32), _c(hyp, np.float32)
    assert inliers.dtype == np.uint8 and inliers.flags.c_contiguous and hyp.shape[-1] == 3
    tn, vn, _ = direct.shape
    hn = hyp.shape[0]
    lib().orc_voting_for_hypothesis_vanishing_point(_p(direct, _f32p), _p(coords, _f32p), _p(hyp, _f32p),
                                                    _p(inliers, _u8p), tn, vn, hn, float(thresh))


def voting_for_hypothesis(direct, coords, hyp, inliers, thresh):
    """In-place on `inliers` (uint8 [hn,vn,tn], pre-zeroed), like the reference."""
    direct, coords, hyp = _c(direct, np.float32), _c(coords, np.float32), _c(hyp, np.float32)
    assert inliers.dtype == np.uint8 and inliers.flags.c_contiguous
    tn, vn, _ = direct.shape
    hn = hyp.shape[0]
    lib().orc_voting_for_hypothesis(_p(direct, _f32p), _p(coords, _f32p), _p(hyp, _f32p), _p(inliers, _u8p),
                                    tn, vn, hn, float(thresh))


def vote_counts(direct, coords, hyp, thresh):
    direct, coords, hyp = _c(direct, np.float32), _c(coords, np.float32), _c(hyp, np.float32)
    tn, vn, _ = direct.shape
    hn = hyp.shape[0]
    counts = np.zeros((hn, vn), np.int64)
    lib().orc_vote_counts(_p(direct, _f32p), _p(coords, _f32p), _p(hyp, _f32p), tn, vn, hn, float(thresh), _p(counts, _i64p))
    return counts


def num_threads() -> int:
    return int(lib().orc_num_threads())


def use_all_host_threads() -> int:
    """Size the OpenMP pool to the cores this process may run on (torchrun sets OMP_NUM_THREADS=1 for its workers)."""
    import os
    n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    lib().orc_set_num_threads(int(n))
    return num_threads()


# ----------------------------------------------------------------------------
# compaction (ransac_voting_gpu.py:533-552 / :339-361)
# ----------------------------------------------------------------------------
def subsample_prob(max_num: int, fg: int) -> np.float32:
    """`max_num / fg.float()` is Tensor.__rtruediv__ = reciprocal(fg) * max_num in fp32."""
    return np.float32(np.float32(1.0) / np.float32(fg)) * np.float32(max_num)


def compact(mask_img, vertex_img, mode: str, min_num: int, max_num: int, selection=None):
    """mask_img [h,w] any int dtype, vertex_img [h,w,vn,2] f32.
    mode 'v3': cur_mask = mask.byte(), fg = sum of byte values (:533-534)
    mode 'eq1': cur_mask = mask == 1                          (:339-340)
    Returns None when the image is skipped (fg < min_num), else
    (coords [tn,2] f32 (x,y), direct [tn,vn,2] f32, flat pixel index [tn] int32)."""
    if mode == "v3":
        cur = np.asarray(mask_img).astype(np.int64).astype(np.uint8)  # wraps mod 256 like .byte()
        fg = int(cur.astype(np.int64).sum())
    else:
        cur = (np.asarray(mask_img) == 1).astype(np.uint8)
        fg = int(cur.sum())
    if fg < min_num:
        return None
    if fg > max_num:
        assert selection is not None, "image needs subsampling: pass `selection` uniforms"
        keep = np.asarray(selection, np.float32) < subsample_prob(max_num, fg)
        cur = cur * keep.astype(np.uint8)
    ys, xs = np.nonzero(cur)  # row-major, like torch.nonzero
    coords = np.stack([xs, ys], 1).astype(np.float32)
    direct = np.ascontiguousarray(np.asarray(vertex_img)[ys, xs], dtype=np.float32)  # [tn,vn,2]
    flat = (ys * cur.shape[1] + xs).astype(np.int32)
    return coords, direct, flat


# ----------------------------------------------------------------------------
# ransac_voting_layer_v3  (ransac_voting_gpu.py:520-604)
# ----------------------------------------------------------------------------
def refine(direct, coords, inl):
    """:585-601 -- weighted normal equations in float64 from float32 terms.
    direct [tn,vn,2], coords [tn,2], inl [vn,tn] {0,1} -> (pts [vn,2] f64, ATA, ATb)."""
    nrm = np.empty_like(direct)
    nrm[:, :, 0] = direct[:, :, 1]
    nrm[:, :, 1] = -direct[:, :, 0]
    nrm = nrm.transpose(1, 0, 2) * inl[:, :, None].astype(np.float32)  # [vn,tn,2] f32
    bb = (nrm * coords[None]).astype(np.float32)  # products in f32 as torch does
    bsum = (bb[:, :, 0] + bb[:, :, 1]).astype(np.float32)  # [vn,tn] two-term f32 sum (:596)
    nb = (nrm * bsum[:, :, None]).astype(np.float32)  # f32 terms of ATb (:598)
    n64 = nrm.astype(np.float64)
    ATA = np.einsum("vti,vtj->vij", n64, n64)
    ATb = nb.astype(np.float64).sum(1)
    vn = direct.shape[1]
    pts = np.zeros((vn, 2))
    for v in range(vn):
        det = ATA[v, 0, 0] * ATA[v, 1, 1] - ATA[v, 0, 1] * ATA[v, 1, 0]
        if det == 0.0 or not np.isfinite(det):
            pts[v] = ATb[v]  # identity fallback (:516-517), per keypoint
        else:
            pts[v] = np.linalg.solve(ATA[v], ATb[v])
    return pts, ATA, ATb


def ransac_voting_layer_v3(mask, vertex, round_hyp_num, inlier_thresh=0.99, confidence=0.99, max_iter=100,
                           min_num=100, max_num=30000, *, idxs=None, selection=None, return_debug=False):
    mask = np.asarray(mask)
    vertex = np.asarray(vertex)
    b, h, w, vn, _ = vertex.shape
    out = np.zeros((b, vn, 2), np.float32)
    dbg = []
    for bi in range(b):
        cp = compact(mask[bi], vertex[bi], "v3", min_num, max_num, None if selection is None else selection[bi])
        if cp is None:
            dbg.append(None)
            continue
        coords, direct, flat = cp
        tn = coords.shape[0]
        cur_idxs = np.asarray(idxs[bi], np.int32)
        assert cur_idxs.shape == (round_hyp_num, vn, 2) and (tn == 0 or cur_idxs.max() < tn)
        hyp = generate_hypothesis(direct, coords, cur_idxs)
        counts = vote_counts(direct, coords, hyp, inlier_thresh)  # [hn,vn]
        win_idx = counts.argmax(0)  # first maximal index, like torch.max(dim=0)
        win_counts = counts[win_idx, np.arange(vn)]
        win_pts = np.zeros((vn, 2), np.float32)
        larger = win_counts > 0  # all_win_ratio(=0) < cur_win_ratio  (:573)
        win_pts[larger] = hyp[win_idx, np.arange(vn)][larger]
        inl = np.zeros((1, vn, tn), np.uint8)
        voting_for_hypothesis(direct, coords, win_pts[None], inl, inlier_thresh)
        pts, ATA, ATb = refine(direct, coords, inl[0])
        out[bi] = pts.astype(np.float32)
        if return_debug:
            dbg.append(dict(coords=coords, direct=direct, flat=flat, hyp=hyp, counts=counts, win_idx=win_idx,
                            win_counts=win_counts, win_pts=win_pts, inlier=inl[0], ATA=ATA, ATb=ATb, pts64=pts))
    return (out, dbg) if return_debug else out


# ----------------------------------------------------------------------------
# ransac_voting_layer_v5  (ransac_voting_gpu.py:769-864)  and  generate_hypothesis  (:989-1040)
# ----------------------------------------------------------------------------
def ransac_voting_layer_v5(mask, vertex, round_hyp_num, inlier_thresh=0.999, confidence=0.99, max_iter=20,
                           min_num=5, max_num=100, *, idxs=None, selection=None, return_debug=False):
    """v3's body with v5's defaults, then the confidence re-vote of the refined (float32) keypoints at the fixed
    threshold 0.999 (:854-856): conf = inlier count / tn in float32; zeros for skipped images (:786-791)."""
    out, dbg = ransac_voting_layer_v3(mask, vertex, round_hyp_num, inlier_thresh, confidence, max_iter, min_num, max_num,
                                      idxs=idxs, selection=selection, return_debug=True)
    b, vn = out.shape[:2]
    conf = np.zeros((b, vn), np.float32)
    for bi in range(b):
        if dbg[bi] is None:
            continue
        tn = dbg[bi]["coords"].shape[0]
        inl = np.zeros((1, vn, tn), np.uint8)
        voting_for_hypothesis(dbg[bi]["direct"], dbg[bi]["coords"], out[bi][None], inl, 0.999)
        conf[bi] = inl[0].astype(np.int64).sum(1).astype(np.float32) / np.float32(tn)
    return (out, conf, dbg) if return_debug else (out, conf)


def generate_hypothesis_layer(mask, vertex, round_hyp_num, inlier_thresh=0.999, confidence=0.99, max_iter=20,
                              min_num=5, max_num=30000, *, idxs=None, selection=None):
    """The wrapper `generate_hypothesis` of ransac_voting_gpu.py:989-1040 (named *_layer here because this module's
    `generate_hypothesis` is the kernel): (hyp [b,hn,vn,2] f32, counts [b,hn,vn] i64).  Skipped images give zeros (the
    reference raises NameError on them, :1009)."""
    mask, vertex = np.asarray(mask), np.asarray(vertex)
    b, h, w, vn, _ = vertex.shape
    hyps = np.zeros((b, round_hyp_num, vn, 2), np.float32)
    counts = np.zeros((b, round_hyp_num, vn), np.int64)
    for bi in range(b):
        cp = compact(mask[bi], vertex[bi], "v3", min_num, max_num, None if selection is None else selection[bi])
        if cp is None:
            continue
        coords, direct, _ = cp
        hyps[bi] = generate_hypothesis(direct, coords, np.asarray(idxs[bi], np.int32))
        counts[bi] = vote_counts(direct, coords, hyps[bi], inlier_thresh)
    return hyps, counts


def ransac_motion_voting(mask, vertex):
    """ransac_voting_gpu.py:966-987: per image the mean over foreground pixels of (vertex + coords); zeros if none."""
    mask, vertex = np.asarray(mask), np.asarray(vertex)
    b, h, w, vn, _ = vertex.shape
    out = np.zeros((b, vn, 2), np.float32)
    for bi in range(b):
        cur = np.asarray(mask[bi]).astype(np.int64).astype(np.uint8)
        ys, xs = np.nonzero(cur)
        if len(xs) < 1:
            continue
        coords = np.stack([xs, ys], 1).astype(np.float32)
        cur_vert = (vertex[bi][ys, xs] + coords[:, None, :]).astype(np.float32)   # fp32 terms as torch forms them
        out[bi] = cur_vert.astype(np.float64).mean(0).astype(np.float32)
    return out


# ----------------------------------------------------------------------------
# estimate_voting_distribution_with_mean  (ransac_voting_gpu.py:333-406)
# ----------------------------------------------------------------------------
def estimate_voting_distribution_with_mean(mask, vertex, mean, round_hyp_num=256, min_hyp_num=4096, topk=128,
                                           inlier_thresh=0.99, min_num=20, max_num=30000, output_hyp=False, *,
                                           idxs=None, selection=None, return_debug=False):
    mask = np.asarray(mask)
    vertex = np.asarray(vertex)
    mean = np.asarray(mean, np.float32)
    b, h, w, vn, _ = vertex.shape
    rounds = int(np.ceil(min_hyp_num / round_hyp_num))
    cov = np.zeros((b, vn, 2, 2), np.float32)
    dbg = []
    for bi in range(b):
        cp = compact(mask[bi], vertex[bi], "eq1", min_num, max_num, None if selection is None else selection[bi])
        if cp is None:
            hyp = np.zeros((min_hyp_num, vn, 2), np.float32)      # :343-347
            ratio = np.ones((min_hyp_num, vn), np.float32)
            counts = None
        else:
            coords, direct, flat = cp
            tn = coords.shape[0]
            hyps, cnts = [], []
            for r in range(rounds):
                cur_idxs = np.asarray(idxs[bi, r], np.int32)
                hp = generate_hypothesis(direct, coords, cur_idxs)
                hyps.append(hp)
                cnts.append(vote_counts(direct, coords, hp, inlier_thresh))
            hyp = np.concatenate(hyps, 0)
            counts = np.concatenate(cnts, 0)
            ratio = counts.astype(np.float32) / np.float32(tn)    # :375, tensor/tensor fp32 division
        thr = (ratio.max(0) - np.float32(0.1)).astype(np.float32)  # :394
        wgt = np.where(ratio < thr[None], np.float32(0), ratio).astype(np.float32)  # :395
        d = (hyp - mean[bi][None]).astype(np.float32)              # :398 [hn,vn,2]
        wd = (d * wgt[:, :, None]).astype(np.float32)              # :399
        c = np.einsum("hvi,hvj->vij", d.astype(np.float64), wd.astype(np.float64))
        c /= (wgt.astype(np.float64).sum(0) + 1e-3)[:, None, None]  # :401
        cov[bi] = c.astype(np.float32)
        if return_debug:
            dbg.append(dict(hyp=hyp, counts=counts, ratio=ratio, cov64=c))
    return (mean, cov, dbg) if return_debug else (mean, cov)
