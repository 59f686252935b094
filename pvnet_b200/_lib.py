"""ctypes binding of libpvvote.so (include/pvvote.h).  No fallback: if the CUDA library is missing this raises."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# PVV_LIB_PATH: load another build of the same library (development A/B builds under tools/_build); never a fallback
LIB_PATH = os.environ.get("PVV_LIB_PATH") or os.path.join(_HERE, "libpvvote.so")

PVV_OK = 0
MASK_U8, MASK_I32, MASK_I64, MASK_F32, MASK_LOGITS_F32 = 0, 1, 2, 3, 4
FG_BYTE, FG_EQ1 = 0, 1
FLAG_POOL_OVERFLOW = 1
STATUS_NAMES = {0: "PVV_OK", -1: "PVV_ERR_INVALID", -2: "PVV_ERR_WORKSPACE", -3: "PVV_ERR_CUDA",
                -4: "PVV_ERR_UNSUPPORTED", -5: "PVV_ERR_POOL_OVERFLOW"}


class Job(C.Structure):
    _fields_ = [
        ("mask", C.c_void_p), ("mask_dtype", C.c_int32), ("fg_rule", C.c_int32), ("mask_strides", C.c_int64 * 3),
        ("vertex", C.c_void_p), ("vertex_strides", C.c_int64 * 5),
        ("b", C.c_int32), ("h", C.c_int32), ("w", C.c_int32), ("vn", C.c_int32), ("hn_total", C.c_int32),
        ("inlier_thresh", C.c_float), ("min_num", C.c_int32), ("max_num", C.c_int32),
        ("idxs", C.c_void_p), ("selection", C.c_void_p), ("seed", C.c_uint64), ("image_id_base", C.c_int64),
        ("pool_capacity", C.c_int64), ("workspace", C.c_void_p), ("workspace_bytes", C.c_size_t),
        ("stream", C.c_void_p), ("seg_classes", C.c_int32), ("seg_class_stride", C.c_int64),
        ("status", C.c_void_p),
    ]


class Debug(C.Structure):
    _fields_ = [
        ("tn", C.c_void_p), ("img_off", C.c_void_p), ("pix_xy", C.c_void_p), ("pix_dir", C.c_void_p),
        ("idxs", C.c_void_p), ("hyp", C.c_void_p), ("counts", C.c_void_p), ("win_idx", C.c_void_p),
        ("win_count", C.c_void_p), ("win_pts", C.c_void_p), ("win_inlier", C.c_void_p), ("normal_eq", C.c_void_p),
        ("status", C.c_void_p), ("pool_capacity", C.c_int64), ("force_exact", C.c_int32), ("amb_count", C.c_void_p),
        ("stage_ms", C.c_void_p), ("k3_variant", C.c_int32), ("hyp_inject", C.c_void_p),
    ]


# every symbol include/pvvote.h declares (tests check the library exports exactly these)
SYMBOLS = [
    "pvv_workspace_bytes", "pvv_default_pool_capacity", "pvv_v3", "pvv_v3_noalloc", "pvv_static_workspace_bytes", "pvv_v5", "pvv_hypotheses", "pvv_motion", "pvv_cov", "pvv_workspace_bytes_host", "pvv_set_host_workers",
    "pvv_v3_host", "pvv_cov_host", "pvv_generate_hypothesis", "pvv_voting_for_hypothesis", "pvv_vote_counts",
    "pvv_generate_hypothesis_vanishing_point", "pvv_voting_for_hypothesis_vanishing_point",
    "pvv_uncertainty_pnp", "pvv_uncertainty_pnp_stop", "pvv_make_idxs", "pvv_make_selection", "pvv_fp32_peak_tflops", "pvv_probe_tflops", "pvv_last_launch_count", "pvv_last_cuda_error",
    "pvv_version",
]

_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -m pvnet_b200.build` (nvcc, sm_100a). "
            "pvnet_b200 has no CPU or PyTorch fallback.")
    L = C.CDLL(LIB_PATH)
    P, I, F, V = C.POINTER, C.c_int, C.c_float, C.c_void_p
    L.pvv_workspace_bytes.argtypes = [P(Job)]
    L.pvv_workspace_bytes.restype = C.c_size_t
    L.pvv_workspace_bytes_host.argtypes = [P(Job)]
    L.pvv_workspace_bytes_host.restype = C.c_size_t
    L.pvv_default_pool_capacity.argtypes = [P(Job)]
    L.pvv_default_pool_capacity.restype = C.c_int64
    L.pvv_v3.argtypes = [P(Job), F, I, V, P(Debug)]
    L.pvv_v5.argtypes = [P(Job), F, I, V, V, P(Debug)]
    L.pvv_v3_noalloc.argtypes = [P(Job), F, I, V]
    L.pvv_static_workspace_bytes.argtypes = []
    L.pvv_static_workspace_bytes.restype = C.c_size_t
    L.pvv_hypotheses.argtypes = [P(Job), V, V, P(Debug)]
    L.pvv_motion.argtypes = [P(Job), V]
    L.pvv_cov.argtypes = [P(Job), V, V, P(Debug)]
    L.pvv_v3_host.argtypes = [P(Job), F, I, V, P(C.c_int64), P(C.c_int64)]
    L.pvv_cov_host.argtypes = [P(Job), V, V, P(C.c_int64), P(C.c_int64)]
    L.pvv_generate_hypothesis.argtypes = [V, V, V, I, I, I, V, V]
    L.pvv_voting_for_hypothesis.argtypes = [V, V, V, V, I, I, I, F, V]
    L.pvv_generate_hypothesis_vanishing_point.argtypes = [V, V, V, I, I, I, V, V]
    L.pvv_voting_for_hypothesis_vanishing_point.argtypes = [V, V, V, V, I, I, I, F, V]
    L.pvv_vote_counts.argtypes = [V, V, V, I, I, I, F, V, V]
    L.pvv_uncertainty_pnp.argtypes = [V, V, I, I, V, C.c_int64, V, C.c_int64, I, I, V, V, V]
    L.pvv_uncertainty_pnp_stop.argtypes = [V, V, I, I, V, C.c_int64, V, C.c_int64, I, I, I, C.c_double, V, V, V]
    L.pvv_make_idxs.argtypes = [C.c_uint64, C.c_int64, I, I, I, V, V, V]
    L.pvv_make_selection.argtypes = [C.c_uint64, C.c_int64, I, I, I, V, V]
    L.pvv_fp32_peak_tflops.argtypes = [I, V]
    L.pvv_fp32_peak_tflops.restype = C.c_double
    L.pvv_probe_tflops.argtypes = [I, I, V]
    L.pvv_probe_tflops.restype = C.c_double
    L.pvv_version.restype = C.c_char_p
    L.pvv_set_host_workers.argtypes = [I]
    # several ranks on one host (torchrun sets LOCAL_WORLD_SIZE): the ranks share the host's hardware threads.  Measured
    # with the AVX-512 narrowing (round 2, `e2e` of bench.py, total images/s, off -> on): 2 ranks x 8 workers 25.3 k -> 37.5 k
    # (6: 35.1 k, 4: 29.2 k, 2: 17.1 k), 4 ranks x 4 workers 31.4 k -> 44.0 k (3: 41.0 k), 8 ranks x 4 workers 51.0 k -> 51.6 k
    # (2: 44.9 k).  So: every rank takes its share of ALL hardware threads, and prepares on the host when that is >= 3.
    # PVV_HOST_WORKERS overrides (0 = off).
    ranks = int(os.environ.get("LOCAL_WORLD_SIZE", "1") or 1)
    forced = os.environ.get("PVV_HOST_WORKERS")
    if forced is not None:
        L.pvv_set_host_workers(int(forced))
    elif ranks > 1:
        per_rank = (os.cpu_count() or 2) // ranks
        L.pvv_set_host_workers(min(per_rank, 16) if per_rank >= 3 else 0)
    _lib = L
    return L


def check(status: int, what: str):
    if status != PVV_OK:
        extra = f" (cudaError {lib().pvv_last_cuda_error()})" if status == -3 else ""
        raise RuntimeError(f"{what} failed: {STATUS_NAMES.get(status, status)}{extra}")
