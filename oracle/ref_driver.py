"""Drive the REFERENCE's own CUDA kernels (oracle/_ref/ref_ransac_voting.so, built by oracle/build.py from
/root/reference/lib/ransac_voting_gpu_layer/src) on the GPU box.  TEST INFRASTRUCTURE ONLY.

The reference's Python wrapper cannot travel to the GPU box and does not run on torch 2.11 anyway,
so the per-image control flow is restated here in torch, following
  ransac_voting_layer_v3                  ransac_voting_gpu.py:520-604
  estimate_voting_distribution_with_mean  ransac_voting_gpu.py:333-406
with the three documented changes: bool masks for masked_select, torch.linalg.inv instead of torch.gesv in b_inv,
and injected `idxs` / `selection` instead of .random_/.uniform_.  One voting round is run (the reference re-votes
identical idxs every round).  Intermediates are returned so tests can compare them bit for bit.
"""
from __future__ import annotations

import importlib.util
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_ref", "ref_ransac_voting.so")
_mod = None


def available() -> bool:
    return os.path.exists(_SO)


def ext():
    """The reference's pybind module, loaded from oracle/_ref."""
    global _mod
    if _mod is None:
        spec = importlib.util.spec_from_file_location("ref_ransac_voting", _SO)
        _mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(_mod)
    return _mod


def _compact(mask_img, vertex_img, rule, min_num, max_num, selection):
    if rule == "v3":
        cur = mask_img.byte()
        fg = torch.sum(cur)
    else:
        cur = mask_img == 1
        fg = torch.sum(cur)
    if fg < min_num:
        return None
    if fg > max_num:
        keep = selection < (max_num / fg.float())
        cur = cur * keep
    cur = cur.bool()
    coords = torch.nonzero(cur).float()[:, [1, 0]].contiguous()
    direct = vertex_img.masked_select(cur[:, :, None, None]).view(coords.shape[0], vertex_img.shape[2], 2).contiguous()
    return coords, direct


def v3(mask, vertex, hn, thresh, min_num, max_num, idxs, selection=None):
    rv = ext()
    b, h, w, vn, _ = vertex.shape
    dev = mask.device
    out = torch.zeros((b, vn, 2), dtype=torch.float32, device=dev)
    dbg = []
    for bi in range(b):
        cp = _compact(mask[bi], vertex[bi], "v3", min_num, max_num, None if selection is None else selection[bi])
        if cp is None:
            dbg.append(None)
            continue
        coords, direct = cp
        tn = coords.shape[0]
        cur_idxs = idxs[bi].to(device=dev, dtype=torch.int32).contiguous()
        hyp = rv.generate_hypothesis(direct, coords, cur_idxs)
        inl = torch.zeros((hn, vn, tn), dtype=torch.uint8, device=dev)
        rv.voting_for_hypothesis(direct, coords, hyp, inl, thresh)
        counts = torch.sum(inl, 2)
        win_counts, win_idx = torch.max(counts, 0)
        win_pts = torch.zeros((vn, 2), dtype=torch.float32, device=dev)
        larger = win_counts > 0
        win_pts[larger] = hyp[win_idx, torch.arange(vn, device=dev)][larger]
        winl = torch.zeros((1, vn, tn), dtype=torch.uint8, device=dev)
        rv.voting_for_hypothesis(direct, coords, win_pts[None].contiguous(), winl, thresh)
        wf = winl[0].float()                                   # [vn,tn]
        normal = torch.stack([direct[:, :, 1], -direct[:, :, 0]], 2).permute(1, 0, 2) * wf[:, :, None]
        bb = torch.sum(normal * coords[None], 2)
        ATA = torch.matmul(normal.permute(0, 2, 1), normal)
        ATb = torch.sum(normal * bb[:, :, None], 1)
        try:
            inv = torch.linalg.inv(ATA)
        except Exception:
            inv = torch.eye(2, device=dev).expand_as(ATA)
        out[bi] = torch.matmul(inv, ATb[:, :, None])[:, :, 0]
        dbg.append(dict(coords=coords, direct=direct, hyp=hyp, counts=counts, win_idx=win_idx, win_pts=win_pts,
                        inlier=winl[0]))
    return out, dbg


def cov(mask, vertex, mean, round_hyp, min_hyp, thresh, min_num, max_num, idxs, selection=None):
    rv = ext()
    b, h, w, vn, _ = vertex.shape
    dev = mask.device
    rounds = -(-min_hyp // round_hyp)
    hyps, ratios = [], []
    for bi in range(b):
        cp = _compact(mask[bi], vertex[bi], "eq1", min_num, max_num, None if selection is None else selection[bi])
        if cp is None:
            hyps.append(torch.zeros((1, min_hyp, vn, 2), device=dev))
            ratios.append(torch.ones((1, min_hyp, vn), device=dev))
            continue
        coords, direct = cp
        tn = coords.shape[0]
        hs, rs = [], []
        for r in range(rounds):
            cur_idxs = idxs[bi, r].to(device=dev, dtype=torch.int32).contiguous()
            hp = rv.generate_hypothesis(direct, coords, cur_idxs)
            inl = torch.zeros((round_hyp, vn, tn), dtype=torch.uint8, device=dev)
            rv.voting_for_hypothesis(direct, coords, hp, inl, thresh)
            hs.append(hp)
            rs.append(torch.sum(inl, 2).float() / torch.tensor(float(tn), device=dev))
        hyps.append(torch.cat(hs, 0)[None])
        ratios.append(torch.cat(rs, 0)[None])
    hyp = torch.cat(hyps, 0).permute(0, 2, 1, 3)
    ratio = torch.cat(ratios, 0).permute(0, 2, 1).clone()
    thr = torch.max(ratio, 2)[0] - 0.1
    ratio[ratio < thr[:, :, None]] = 0.0
    diff = hyp - mean[:, :, None]
    c = torch.matmul(diff.transpose(2, 3), diff * ratio[..., None])
    c = c / (torch.sum(ratio, 2)[:, :, None, None] + 1e-3)
    return mean, c
