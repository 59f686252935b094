"""Drop-in mirror of the reference module lib/ransac_voting_gpu_layer/ransac_voting_gpu.py for the voting hot path.

Same names, argument meaning, defaults, tensor layouts and output shapes as the reference:
  ransac_voting_layer_v3                  (reference ransac_voting_gpu.py:520-604)
  estimate_voting_distribution_with_mean  (reference ransac_voting_gpu.py:333-406)
  ransac_voting.generate_hypothesis / voting_for_hypothesis   (reference src/ransac_voting.cpp:102-107)
but the whole batch runs as one fused sequence of sm_100a kernels (libpvvote.so, include/pvvote.h) on the current
CUDA stream with no host synchronisation.  PyTorch only supplies device memory and the stream.

Extra keyword-only arguments (all optional): `idxs`, `selection` inject the random draws (parity runs), `seed` /
`image_id_base` key the library's own counter-based draw stream, `return_debug` also returns intermediate tensors.
CPU tensors take the host-buffer entry points (pvv_v3_host / pvv_cov_host) and return CPU tensors.
There is no fallback: without libpvvote.so or without a CUDA device every call raises.
"""
from __future__ import annotations

import ctypes as C
import math
import os

import torch

from . import _lib

__all__ = ["ransac_voting_layer_v3", "ransac_voting_layer_v5", "generate_hypothesis", "ransac_motion_voting", "estimate_voting_distribution_with_mean", "ransac_voting", "torch_extension", "make_idxs",
           "make_selection", "fp32_peak_tflops", "last_launch_count"]

_MASK_DT = {torch.uint8: _lib.MASK_U8, torch.bool: _lib.MASK_U8, torch.int32: _lib.MASK_I32,
            torch.int64: _lib.MASK_I64, torch.float32: _lib.MASK_F32}


def _stream_ptr(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


_SEED_GEN = None


def _fresh_seed():
    """Like the reference's .random_(): a new draw stream per call.  A private generator, seeded once from torch's global
    seed, so that torch.manual_seed() still makes runs reproducible but no call advances the user's own RNG state."""
    global _SEED_GEN
    if _SEED_GEN is None:
        _SEED_GEN = torch.Generator(device="cpu")
        _SEED_GEN.manual_seed(torch.initial_seed() & (2 ** 63 - 1))
    return int(torch.randint(0, 2 ** 62, (1,), generator=_SEED_GEN).item())


# tuning / test knob: PVV_K3_VARIANT=<n> forces inlier-counting kernel variant n (read once, at import; the C ABI takes
# it through pvv_debug.k3_variant -- the library itself never reads the environment)
_K3_VARIANT_ENV = os.environ.get("PVV_K3_VARIANT")


def _prep_mask(mask):
    if mask.dim() == 4:  # seg_pred logits [b,C,h,w]: torch.argmax(seg_pred, 1) is fused into the compaction pass
        return mask if mask.dtype == torch.float32 else mask.to(torch.float32)
    if mask.dtype not in _MASK_DT:
        mask = mask.to(torch.float32) if mask.dtype.is_floating_point else mask.to(torch.int32)
    return mask


def _make_job(mask, vertex, hn_total, thresh, min_num, max_num, fg_rule, idxs, selection, seed, image_id_base,
              pool_capacity):
    b, h, w, vn, two = vertex.shape
    logits = mask.dim() == 4
    if two != 2 or (mask.shape[0], mask.shape[-2], mask.shape[-1]) != (b, h, w) or mask.dim() not in (3, 4):
        raise ValueError(f"expected mask [b,h,w] (or seg_pred [b,C,h,w]) and vertex [b,h,w,vn,2], got "
                         f"{tuple(mask.shape)} / {tuple(vertex.shape)}")
    if vertex.dtype != torch.float32:
        raise TypeError("vertex must be float32")
    j = _lib.Job()
    j.mask, j.fg_rule = mask.data_ptr(), fg_rule
    if logits:
        j.mask_dtype = _lib.MASK_LOGITS_F32
        sb, sc, sh, sw = mask.stride()
        j.mask_strides = (C.c_int64 * 3)(sb, sh, sw)
        j.seg_classes, j.seg_class_stride = mask.shape[1], sc
    else:
        j.mask_dtype = _MASK_DT[mask.dtype]
        j.mask_strides = (C.c_int64 * 3)(*mask.stride())
    j.vertex = vertex.data_ptr()
    j.vertex_strides = (C.c_int64 * 5)(*vertex.stride())
    j.b, j.h, j.w, j.vn, j.hn_total = b, h, w, vn, int(hn_total)
    j.inlier_thresh, j.min_num, j.max_num = float(thresh), int(min_num), int(min(max_num, 2 ** 31 - 1))
    j.idxs = idxs.data_ptr() if idxs is not None else None
    j.selection = selection.data_ptr() if selection is not None else None
    j.seed = int(seed) & (2 ** 64 - 1)
    j.image_id_base = int(image_id_base)
    j.pool_capacity = int(pool_capacity or 0)
    return j


def _check_draws(idxs, selection, b, h, w, vn, hn_total, device):
    if idxs is not None:
        idxs = idxs.to(device=device, dtype=torch.int32).reshape(b, hn_total, vn, 2).contiguous()
    if selection is not None:
        selection = selection.to(device=device, dtype=torch.float32).reshape(b, h, w).contiguous()
    return idxs, selection


def _alloc_debug(job, dev, b, vn, hn, want_pool):
    cap = int(_lib.lib().pvv_default_pool_capacity(C.byref(job)))
    t = {
        "tn": torch.empty(b, dtype=torch.int32, device=dev),
        "img_off": torch.empty(b, dtype=torch.int64, device=dev),
        "idxs": torch.empty((b, hn, vn, 2), dtype=torch.int32, device=dev),
        "hyp": torch.empty((b, hn, vn, 2), dtype=torch.float32, device=dev),
        "counts": torch.empty((b, hn, vn), dtype=torch.int32, device=dev),
        "win_idx": torch.zeros((b, vn), dtype=torch.int32, device=dev),
        "win_count": torch.zeros((b, vn), dtype=torch.int32, device=dev),
        "win_pts": torch.zeros((b, vn, 2), dtype=torch.float32, device=dev),
        "normal_eq": torch.zeros((b, vn, 5), dtype=torch.float64, device=dev),
        "status": torch.zeros(1, dtype=torch.int32, device=dev),
        "amb_count": torch.zeros(1, dtype=torch.int64, device=dev),
    }
    if want_pool:
        t["pix_xy"] = torch.empty(cap, dtype=torch.int32, device=dev)
        t["pix_dir"] = torch.empty((vn, cap, 2), dtype=torch.float32, device=dev)
        t["win_inlier"] = torch.zeros((vn, cap), dtype=torch.uint8, device=dev)
    d = _lib.Debug()
    for k, v in t.items():
        setattr(d, k, v.data_ptr())
    d.pool_capacity = cap
    return d, t


def _run(kind, mask, vertex, mean, hn_total, thresh, confidence, max_iter, min_num, max_num, idxs, selection, seed,
         image_id_base, return_debug, force_exact, pool_capacity, stage_ms=None, k3_variant=None, hyp_inject=None):
    L = _lib.lib()
    b, h, w, vn, _ = vertex.shape
    if k3_variant is None and _K3_VARIANT_ENV is not None:
        k3_variant = int(_K3_VARIANT_ENV)
    fg_rule = _lib.FG_EQ1 if kind == "cov" else _lib.FG_BYTE
    mask = _prep_mask(mask)
    if seed is None:
        seed = _fresh_seed()
    host = not mask.is_cuda
    if host != (not vertex.is_cuda):
        raise ValueError("mask and vertex must live on the same side (both CUDA or both CPU)")
    if not host and mask.device != vertex.device:
        raise ValueError(f"mask is on {mask.device} but vertex is on {vertex.device}")
    if host:
        for name, t in (("idxs", idxs), ("selection", selection), ("mean", mean)):
            if t is not None and t.is_cuda:
                raise ValueError(f"host-buffer call: `{name}` must be a CPU tensor like mask and vertex")
    if not torch.cuda.is_available():
        raise RuntimeError("pvnet_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    dev = torch.device("cuda", torch.cuda.current_device()) if host else mask.device
    out_shape = (b, vn, 2, 2) if kind == "cov" else (b, vn, 2)

    if host:
        if return_debug:
            raise ValueError("return_debug is only available for CUDA inputs")
        if kind not in ("v3", "cov"):
            raise ValueError("host-buffer calls are available for ransac_voting_layer_v3 and "
                             "estimate_voting_distribution_with_mean")
        idxs_h = None if idxs is None else idxs.to(dtype=torch.int32).reshape(b, hn_total, vn, 2).contiguous()
        sel_h = None if selection is None else selection.to(dtype=torch.float32).reshape(b, h, w).contiguous()
        job = _make_job(mask, vertex, hn_total, thresh, min_num, max_num, fg_rule, idxs_h, sel_h, seed, image_id_base,
                        pool_capacity)
        with torch.cuda.device(dev):
            nbytes = L.pvv_workspace_bytes_host(C.byref(job))
            ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
            job.workspace, job.workspace_bytes, job.stream = ws.data_ptr(), nbytes, _stream_ptr(dev)
            out = torch.empty(out_shape, dtype=torch.float32, pin_memory=True)
            h2d, d2h = C.c_int64(0), C.c_int64(0)
            if kind == "v3":
                st = L.pvv_v3_host(C.byref(job), float(confidence), int(max_iter), out.data_ptr(), C.byref(h2d), C.byref(d2h))
            else:
                mean_h = mean.to(dtype=torch.float32).contiguous()
                st = L.pvv_cov_host(C.byref(job), mean_h.data_ptr(), out.data_ptr(), C.byref(h2d), C.byref(d2h))
        _lib.check(st, f"pvv_{kind}_host")
        out.transfer_bytes = (h2d.value, d2h.value)
        return out

    with torch.cuda.device(dev):
        idxs, selection = _check_draws(idxs, selection, b, h, w, vn, hn_total, dev)
        job = _make_job(mask, vertex, hn_total, thresh, min_num, max_num, fg_rule, idxs, selection, seed, image_id_base,
                        pool_capacity)
        nbytes = L.pvv_workspace_bytes(C.byref(job))
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        job.workspace, job.workspace_bytes, job.stream = ws.data_ptr(), nbytes, _stream_ptr(dev)
        out = torch.empty(out_shape, dtype=torch.float32, device=dev)
        dbg_ref, dbg_t = None, None
        if return_debug or force_exact:
            d, dbg_t = _alloc_debug(job, dev, b, vn, hn_total, want_pool=return_debug)
            d.force_exact = 1 if force_exact else 0
            dbg_ref = C.byref(d)
        elif stage_ms is not None or k3_variant is not None or hyp_inject is not None:
            d = _lib.Debug()
            dbg_ref = C.byref(d)
        if k3_variant is not None:
            d.k3_variant = int(k3_variant) + 1
        if hyp_inject is not None:
            hyp_inject = hyp_inject.to(device=dev, dtype=torch.float32).reshape(b, hn_total, vn, 2).contiguous()
            d.hyp_inject = hyp_inject.data_ptr()
        # A pool overflow can only happen with a user-set pool_capacity or an injected `selection`; those calls check the
        # status word (one 4-byte read, which synchronises) and raise instead of returning zeroed images silently.
        status = None
        if pool_capacity or selection is not None:
            status = torch.zeros(1, dtype=torch.int32, device=dev)
            job.status = status.data_ptr()
        if stage_ms is not None:
            buf = (C.c_float * 5)()
            d.stage_ms = C.cast(buf, C.c_void_p)
        extra = None
        if kind == "v3":
            st = L.pvv_v3(C.byref(job), float(confidence), int(max_iter), out.data_ptr(), dbg_ref)
        elif kind == "v5":
            extra = torch.empty((b, vn), dtype=torch.float32, device=dev)
            st = L.pvv_v5(C.byref(job), float(confidence), int(max_iter), out.data_ptr(), extra.data_ptr(), dbg_ref)
        elif kind == "hyp":
            out = torch.empty((b, hn_total, vn, 2), dtype=torch.float32, device=dev)
            extra = torch.empty((b, hn_total, vn), dtype=torch.int32, device=dev)
            st = L.pvv_hypotheses(C.byref(job), out.data_ptr(), extra.data_ptr(), dbg_ref)
        else:
            mean_d = mean.to(device=dev, dtype=torch.float32).contiguous()
            st = L.pvv_cov(C.byref(job), mean_d.data_ptr(), out.data_ptr(), dbg_ref)
        _lib.check(st, f"pvv_{kind}")
        ws.record_stream(torch.cuda.current_stream(dev))
        if status is not None and not return_debug and not torch.cuda.is_current_stream_capturing():
            if int(status.item()) & _lib.FLAG_POOL_OVERFLOW:
                raise RuntimeError(f"pvv_{kind}: pool overflow -- the compacted foreground pixels do not fit pool_capacity="
                                   f"{pool_capacity or 'default'}; images that did not fit were dropped (raise pool_capacity)")
        if stage_ms is not None:
            stage_ms.append(dict(zip(("compact", "hypotheses", "count", "finalize", "total"), list(buf))))
    if extra is not None:
        out = (out, extra)
    if return_debug:
        dbg_t["seed"] = seed
        return out, dbg_t
    return out


def ransac_voting_layer_v3(mask, vertex, round_hyp_num, inlier_thresh=0.99, confidence=0.99, max_iter=100,
                           min_num=100, max_num=30000, *, idxs=None, selection=None, seed=None, image_id_base=0,
                           return_debug=False, force_exact=False, pool_capacity=None, stage_ms=None, k3_variant=None,
                           hyp_inject=None):
    """mask [b,h,w] (any integer dtype), vertex [b,h,w,vn,2] float32 (any strides) -> keypoints [b,vn,2] float32.

    Reference: ransac_voting_gpu.py:520-604.  Images with fewer than `min_num` foreground pixels give zeros; images
    with more than `max_num` are Bernoulli-subsampled.  `confidence` / `max_iter` are accepted but cannot change the
    result: the reference draws idxs once (:553) and re-votes the same hypotheses every round."""
    return _run("v3", mask, vertex, None, round_hyp_num, inlier_thresh, confidence, max_iter, min_num, max_num, idxs,
                selection, seed, image_id_base, return_debug, force_exact, pool_capacity, stage_ms, k3_variant, hyp_inject)


def ransac_voting_layer_v5(mask, vertex, round_hyp_num, inlier_thresh=0.999, confidence=0.99, max_iter=20, min_num=5,
                           max_num=100, *, idxs=None, selection=None, seed=None, image_id_base=0, return_debug=False,
                           force_exact=False, pool_capacity=None, k3_variant=None):
    """-> (keypoints [b,vn,2], confidence [b,vn]).  Reference: ransac_voting_gpu.py:769-864 = v3 with its own defaults
    plus a re-vote of the refined keypoints at the fixed threshold 0.999 (:854-856); confidence = inlier count / tn,
    zeros for skipped images (:786-791)."""
    r = _run("v5", mask, vertex, None, round_hyp_num, inlier_thresh, confidence, max_iter, min_num, max_num, idxs,
             selection, seed, image_id_base, return_debug, force_exact, pool_capacity, None, k3_variant)
    if return_debug:
        return r[0][0], r[0][1], r[1]
    return r


def generate_hypothesis(mask, vertex, round_hyp_num, inlier_thresh=0.999, confidence=0.99, max_iter=20, min_num=5,
                        max_num=30000, *, idxs=None, selection=None, seed=None, image_id_base=0, return_debug=False,
                        force_exact=False, pool_capacity=None, k3_variant=None, hyp_inject=None):
    """-> (hypotheses [b,hn,vn,2] float32, inlier counts [b,hn,vn] int64).  Reference: ransac_voting_gpu.py:989-1040 (what
    tools/demo.py:127 feeds to visualize_hypothesis).  Skipped images give zeros (the reference raises NameError there:
    `batch_win_pts` is undefined at :1009)."""
    r = _run("hyp", mask, vertex, None, round_hyp_num, inlier_thresh, confidence, max_iter, min_num, max_num, idxs,
             selection, seed, image_id_base, return_debug, force_exact, pool_capacity, None, k3_variant, hyp_inject)
    if return_debug:
        return r[0][0], r[0][1].long(), r[1]
    return r[0], r[1].long()


def ransac_motion_voting(mask, vertex):
    """-> [b,vn,2]: mean over the foreground pixels of (vertex + pixel coordinate); zeros without foreground.
    Reference: ransac_voting_gpu.py:966-987 (MotionEvalWrapper, tools/train_linemod.py:127-133)."""
    if not torch.cuda.is_available() or not mask.is_cuda or not vertex.is_cuda:
        raise RuntimeError("ransac_motion_voting needs CUDA tensors (sm_100a); there is no CPU fallback")
    L = _lib.lib()
    b, h, w, vn, _ = vertex.shape
    mask = _prep_mask(mask)
    dev = mask.device
    with torch.cuda.device(dev):
        job = _make_job(mask, vertex, 1, 0.99, 1, 2 ** 31 - 1, _lib.FG_BYTE, None, None, 0, 0, b * h * w)
        nbytes = L.pvv_workspace_bytes(C.byref(job))
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        job.workspace, job.workspace_bytes, job.stream = ws.data_ptr(), nbytes, _stream_ptr(dev)
        out = torch.empty((b, vn, 2), dtype=torch.float32, device=dev)
        _lib.check(L.pvv_motion(C.byref(job), out.data_ptr()), "pvv_motion")
        ws.record_stream(torch.cuda.current_stream(dev))
    return out


def estimate_voting_distribution_with_mean(mask, vertex, mean, round_hyp_num=256, min_hyp_num=4096, topk=128,
                                           inlier_thresh=0.99, min_num=20, max_num=30000, output_hyp=False, *,
                                           idxs=None, selection=None, seed=None, image_id_base=0, return_debug=False,
                                           force_exact=False, pool_capacity=None, stage_ms=None, k3_variant=None):
    """-> (mean, cov [b,vn,2,2]).  Reference: ransac_voting_gpu.py:333-406 (`topk`, `output_hyp` are ignored there too).
    `idxs`, when injected, is int32 [b, rounds, round_hyp_num, vn, 2] with rounds = ceil(min_hyp_num/round_hyp_num)."""
    rounds = int(math.ceil(min_hyp_num / round_hyp_num))
    r = _run("cov", mask, vertex, mean, rounds * round_hyp_num, inlier_thresh, 0.0, 0, min_num, max_num, idxs, selection,
             seed, image_id_base, return_debug, force_exact, pool_capacity, stage_ms, k3_variant)
    if return_debug:
        return mean, r[0], r[1]
    return mean, r


# ------------------------------------------------------------------------------------------------
# primitive-compatible module object: `ransac_voting.generate_hypothesis(direct, coords, idxs)` etc.
# ------------------------------------------------------------------------------------------------
def _chk(t, name, dtype):
    # same errors as CHECK_INPUT in the reference shim (ransac_voting.cpp:7-9)
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")
    if not t.is_contiguous():
        raise RuntimeError(f"{name} must be contiguous")
    if t.dtype != dtype:
        raise RuntimeError(f"{name} must be {dtype}")


class _RansacVoting:
    """Stands in for the reference's pybind module `ransac_voting` (src/ransac_voting.cpp:102-107)."""

    @staticmethod
    def generate_hypothesis(direct, coords, idxs):
        _chk(direct, "direct", torch.float32); _chk(coords, "coords", torch.float32); _chk(idxs, "idxs", torch.int32)
        tn, vn, _ = direct.shape
        hn = idxs.shape[0]
        hyp = torch.empty((hn, vn, 2), dtype=torch.float32, device=direct.device)
        with torch.cuda.device(direct.device):
            st = _lib.lib().pvv_generate_hypothesis(direct.data_ptr(), coords.data_ptr(), idxs.data_ptr(), tn, vn, hn,
                                                    hyp.data_ptr(), _stream_ptr(direct.device))
        _lib.check(st, "pvv_generate_hypothesis")
        return hyp

    @staticmethod
    def voting_for_hypothesis(direct, coords, hypo_pts, inliers, inlier_thresh):
        _chk(direct, "direct", torch.float32); _chk(coords, "coords", torch.float32)
        _chk(hypo_pts, "hypo_pts", torch.float32); _chk(inliers, "inliers", torch.uint8)
        tn, vn, _ = direct.shape
        hn = hypo_pts.shape[0]
        with torch.cuda.device(direct.device):
            st = _lib.lib().pvv_voting_for_hypothesis(direct.data_ptr(), coords.data_ptr(), hypo_pts.data_ptr(),
                                                      inliers.data_ptr(), tn, vn, hn, float(inlier_thresh),
                                                      _stream_ptr(direct.device))
        _lib.check(st, "pvv_voting_for_hypothesis")

    @staticmethod
    def generate_hypothesis_vanishing_point(direct, coords, idxs):
        """src/ransac_voting.cpp:64-75: homogeneous hypotheses [hn,vn,3]."""
        _chk(direct, "direct", torch.float32); _chk(coords, "coords", torch.float32); _chk(idxs, "idxs", torch.int32)
        tn, vn, _ = direct.shape
        hn = idxs.shape[0]
        hyp = torch.empty((hn, vn, 3), dtype=torch.float32, device=direct.device)
        with torch.cuda.device(direct.device):
            st = _lib.lib().pvv_generate_hypothesis_vanishing_point(direct.data_ptr(), coords.data_ptr(), idxs.data_ptr(),
                                                                    tn, vn, hn, hyp.data_ptr(), _stream_ptr(direct.device))
        _lib.check(st, "pvv_generate_hypothesis_vanishing_point")
        return hyp

    @staticmethod
    def voting_for_hypothesis_vanishing_point(direct, coords, hypo_pts, inliers, inlier_thresh):
        """src/ransac_voting.cpp:85-98: in place on inliers [hn,vn,tn] u8."""
        _chk(direct, "direct", torch.float32); _chk(coords, "coords", torch.float32)
        _chk(hypo_pts, "hypo_pts", torch.float32); _chk(inliers, "inliers", torch.uint8)
        tn, vn, _ = direct.shape
        hn = hypo_pts.shape[0]
        with torch.cuda.device(direct.device):
            st = _lib.lib().pvv_voting_for_hypothesis_vanishing_point(direct.data_ptr(), coords.data_ptr(),
                                                                      hypo_pts.data_ptr(), inliers.data_ptr(), tn, vn, hn,
                                                                      float(inlier_thresh), _stream_ptr(direct.device))
        _lib.check(st, "pvv_voting_for_hypothesis_vanishing_point")

    @staticmethod
    def vote_counts(direct, coords, hypo_pts, inlier_thresh):
        _chk(direct, "direct", torch.float32); _chk(coords, "coords", torch.float32)
        _chk(hypo_pts, "hypo_pts", torch.float32)
        tn, vn, _ = direct.shape
        hn = hypo_pts.shape[0]
        counts = torch.zeros((hn, vn), dtype=torch.int32, device=direct.device)
        with torch.cuda.device(direct.device):
            st = _lib.lib().pvv_vote_counts(direct.data_ptr(), coords.data_ptr(), hypo_pts.data_ptr(), tn, vn, hn,
                                            float(inlier_thresh), counts.data_ptr(), _stream_ptr(direct.device))
        _lib.check(st, "pvv_vote_counts")
        return counts


ransac_voting = _RansacVoting()

_EXT = None


def torch_extension():
    """The PyTorch C++ extension that replaces the reference's pybind module `ransac_voting`
    (src/ransac_voting.cpp:102-107): pvnet_b200/ransac_voting.so, built by `python -m pvnet_b200.build` from
    csrc/ransac_voting_shim.cpp -- a thin shim over the same C ABI.  Raises if it has not been built."""
    global _EXT
    if _EXT is None:
        import importlib.util
        import os
        path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ransac_voting.so")
        if not os.path.exists(path):
            raise RuntimeError(f"{path} is missing: build it with `python -m pvnet_b200.build`")
        _lib.lib()  # libpvvote.so first, so that the shim's DT_NEEDED entry resolves to the in-tree copy
        spec = importlib.util.spec_from_file_location("ransac_voting", path)
        _EXT = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(_EXT)
    return _EXT


def make_idxs(seed, tn, hn_total, vn, image_id_base=0):
    """Materialise the library's pixel-pair stream: tn int32 [b] (CUDA) -> idxs int32 [b,hn_total,vn,2]."""
    tn = tn.to(torch.int32).contiguous()
    b = tn.shape[0]
    idxs = torch.empty((b, hn_total, vn, 2), dtype=torch.int32, device=tn.device)
    with torch.cuda.device(tn.device):
        st = _lib.lib().pvv_make_idxs(int(seed) & (2 ** 64 - 1), int(image_id_base), b, hn_total, vn, tn.data_ptr(),
                                      idxs.data_ptr(), _stream_ptr(tn.device))
    _lib.check(st, "pvv_make_idxs")
    return idxs


def make_selection(seed, b, h, w, device, image_id_base=0):
    sel = torch.empty((b, h, w), dtype=torch.float32, device=device)
    with torch.cuda.device(device):
        st = _lib.lib().pvv_make_selection(int(seed) & (2 ** 64 - 1), int(image_id_base), b, h, w, sel.data_ptr(),
                                           _stream_ptr(device))
    _lib.check(st, "pvv_make_selection")
    return sel


def fp32_peak_tflops(iters=4096, device=None):
    device = torch.device("cuda", torch.cuda.current_device()) if device is None else device
    with torch.cuda.device(device):
        return float(_lib.lib().pvv_fp32_peak_tflops(int(iters), _stream_ptr(device)))


def last_launch_count():
    return int(_lib.lib().pvv_last_launch_count())
