/*
 * pvvote.h -- C ABI of the B200-native RANSAC keypoint-voting library (libpvvote.so).
 *
 * This is the drop-in boundary for PVNet's voting hot path.  It replaces what the
 * reference binds through pybind11 in
 *     lib/ransac_voting_gpu_layer/src/ransac_voting.cpp:102-107   (module `ransac_voting`)
 * and the per-image Python orchestration around it in
 *     lib/ransac_voting_gpu_layer/ransac_voting_gpu.py:520-604    (ransac_voting_layer_v3)
 *     lib/ransac_voting_gpu_layer/ransac_voting_gpu.py:333-406    (estimate_voting_distribution_with_mean)
 * (paths relative to the reference checkout).
 *
 * Conventions
 *   - plain pointers and sizes only; no torch / ATen types cross this boundary;
 *   - every entry point returns 0 (PVV_OK) or a negative pvv_status; nothing exits the process or throws
 *     (the reference's gpuErrchk calls exit(), src/cuda_common.h:19-26);
 *   - all work is enqueued on the caller's stream (`stream` is a cudaStream_t passed as void*), the current
 *     device is the caller's; no host synchronisation happens inside the device-buffer entry points;
 *   - the caller owns every buffer, including the scratch `workspace` sized by pvv_workspace_bytes();
 *   - re-entrant and thread-safe: no global mutable state (DataParallel calls the layer from one thread
 *     per GPU, tools/train_linemod.py:222-223).
 */
#ifndef PVVOTE_H_
#define PVVOTE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum pvv_status {
    PVV_OK = 0,
    PVV_ERR_INVALID = -1,     /* bad argument (null pointer, non-positive size, unsupported dtype, thresh range) */
    PVV_ERR_WORKSPACE = -2,   /* workspace too small for pvv_workspace_bytes() of this job */
    PVV_ERR_CUDA = -3,        /* a CUDA runtime call failed; pvv_last_cuda_error() has the code */
    PVV_ERR_UNSUPPORTED = -4, /* shape outside compiled limits (vn > 64, w or h > 65535) */
    PVV_ERR_POOL_OVERFLOW = -5 /* host-buffer entry points only (they synchronise): the compacted pixels did not fit
                                  pool_capacity; the outputs are written, with zeros for the images that were dropped */
} pvv_status;

/* mask element types accepted (the reference takes whatever `.byte()` / `== 1` accept) */
enum {
    PVV_MASK_U8 = 0, PVV_MASK_I32 = 1, PVV_MASK_I64 = 2, PVV_MASK_F32 = 3,
    /* `mask` is the segmentation head's output seg_pred [b,C,h,w] float32; the label of a pixel is
     * torch.argmax(seg_pred, 1) (first maximal class, a NaN counts as maximal) computed on the fly by the compaction
     * pass, so the int64 mask of EvalWrapper.forward (tools/demo.py:46-55, tools/train_linemod.py:254) is never
     * written.  mask_strides = strides of (b,h,w); the class stride and count are seg_class_stride / seg_classes.
     * Device-buffer entry points only. */
    PVV_MASK_LOGITS_F32 = 4
};

/* which foreground rule to apply */
enum {
    PVV_FG_BYTE = 0, /* ransac_voting_layer_v3:   fg = (uint8)mask != 0, count = sum of byte values (:533-534) */
    PVV_FG_EQ1 = 1   /* estimate_voting_distribution_with_mean: fg = (mask == 1)                     (:339-340) */
};

/* bit flags written to pvv_job.status / pvv_debug.status (device side) */
enum {
    PVV_FLAG_POOL_OVERFLOW = 1 /* compacted pixels exceeded the pool capacity: images are admitted in batch order, one
                                  that does not fit is dropped (zero keypoints) and takes no room, later ones that fit
                                  are kept.  Cannot happen with the default capacity and library-drawn subsampling. */
};

/*
 * One voting job over a batch.  Layouts are the reference's:
 *   mask    [b,h,w]      any PVV_MASK_* type, strides in ELEMENTS (mask_strides[3]); or seg_pred logits [b,C,h,w]
 *           with PVV_MASK_LOGITS_F32
 *   vertex  [b,h,w,vn,2] float32, strides in ELEMENTS (vertex_strides[5]); the network emits this as a
 *           permuted view of [b,2+2vn,h,w] (tools/train_linemod.py:118-120) -- it is consumed in place.
 * Random draws are either injected (parity runs) or produced by the library's counter-based generator
 * keyed by (seed, image_id_base + image, hypothesis, keypoint):
 *   idxs      int32 [b,hn_total,vn,2], each in [0,tn_image)  or NULL
 *   selection float32 [b,h,w] uniforms in [0,1) used only for images with fg > max_num, or NULL
 */
typedef struct pvv_job {
    const void *mask;
    int32_t mask_dtype;
    int32_t fg_rule;
    int64_t mask_strides[3];
    const float *vertex;
    int64_t vertex_strides[5];
    int32_t b, h, w, vn;
    int32_t hn_total;         /* hypotheses per image and keypoint (v3: round_hyp_num; cov: rounds*round_hyp_num) */
    float inlier_thresh;
    int32_t min_num, max_num;
    const int32_t *idxs;      /* device, optional */
    const float *selection;   /* device, optional */
    uint64_t seed;
    int64_t image_id_base;    /* global id of image 0 of this shard (keeps draws shard-invariant) */
    int64_t pool_capacity;    /* max compacted pixels over the whole batch; 0 = library default bound */
    void *workspace;
    size_t workspace_bytes;
    void *stream;
    int32_t seg_classes;      /* PVV_MASK_LOGITS_F32 only: C */
    int64_t seg_class_stride; /* PVV_MASK_LOGITS_F32 only: stride of the class dimension, in elements */
    int32_t *status;          /* device, optional: receives the PVV_FLAG_* bits of this call (0 = clean) on job->stream,
                                 without a host synchronisation -- the caller reads it at its next sync point */
} pvv_job;

/* Optional device-side introspection buffers (any may be NULL); used by the parity tests. */
typedef struct pvv_debug {
    int32_t *tn;          /* [b]   compacted pixel count per image (0 for skipped images) */
    int64_t *img_off;     /* [b]   offset of each image's pixels in the pool */
    uint32_t *pix_xy;     /* [pool] x | y<<16 in row-major compaction order */
    float *pix_dir;       /* [vn][pool][2] raw direction vectors */
    int32_t *idxs;        /* [b,hn_total,vn,2] the pixel-pair indices actually used */
    float *hyp;           /* [b,hn_total,vn,2] hypotheses, reference layout */
    int32_t *counts;      /* [b,hn_total,vn]   inlier counts, reference layout */
    int32_t *win_idx;     /* [b,vn] argmax (first maximal index) */
    int32_t *win_count;   /* [b,vn] */
    float *win_pts;       /* [b,vn,2] winning hypotheses before refinement */
    uint8_t *win_inlier;  /* [vn][pool] winner re-vote inlier flags */
    double *normal_eq;    /* [b,vn,5] ATA00, ATA01, ATA11, ATb0, ATb1 */
    int32_t *status;      /* [1] PVV_FLAG_* */
    int64_t pool_capacity; /* capacity of the pool-sized arrays above, in pixels */
    int32_t force_exact;  /* 1: disable the guard-banded fast filter, evaluate every test exactly */
    uint64_t *amb_count;  /* [1] number of tests that took the exact path */
    float *stage_ms;      /* HOST pointer [5] or NULL: device time of compaction, hypotheses, counting, finalise and the
                             whole call, from CUDA events on job->stream.  Setting it makes the call synchronise. */
    int32_t k3_variant;   /* 0: the library picks the inlier-counting kernel (FP32-pipe up to 128 hypotheses, tensor pipe
                             above); n > 0: force variant n - 1 (tests and tuning: 0..6 and 11..13 FP32-pipe CTA shapes,
                             7 / 8 tensor pipe, 9 tensor pipe with 3 producer warps, 10 the experimental queue-and-defer
                             pair).  Results are bit-identical for every value. */
    const float *hyp_inject; /* device [b,hn_total,vn,2] or NULL: use these hypotheses instead of intersecting pixel pairs
                             (tests only: places hypotheses on the decision cone, tests/test_gpu_parity.py) */
} pvv_debug;

/* Scratch size for a job (depends on b,h,w,vn,hn_total,max_num,pool_capacity only). */
size_t pvv_workspace_bytes(const pvv_job *job);
/* Default pool capacity the library would pick for this job (pixels). */
int64_t pvv_default_pool_capacity(const pvv_job *job);

/*
 * ransac_voting_layer_v3 (ransac_voting_gpu.py:520-604): out_kp [b,vn,2] float32 (device).
 * `confidence` and `max_iter` are accepted for signature fidelity; in the reference every round re-votes the
 * same hypotheses (idxs drawn once at :553), so the result is a function of the first round only.
 */
int pvv_v3(const pvv_job *job, float confidence, int max_iter, float *out_kp, const pvv_debug *dbg);

/*
 * pvv_v3 without any device allocation by the caller or the library.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same inputs, outputs, arithmetic and status codes as pvv_v3, but job->workspace /
 * job->workspace_bytes are IGNORED: the scratch lives in a file-scope static __device__ array of
 * pvv_static_workspace_bytes() = 32 MiB inside the library (every GPU that loads libpvvote.so holds it until the process
 * exits).  A job whose pvv_workspace_bytes() exceeds that is rejected with PVV_ERR_WORKSPACE and nothing is launched
 * (32 MiB covers e.g. 2 images of 480x640, 9 keypoints, 512 hypotheses; check with pvv_workspace_bytes()).
 * The call makes no cudaMalloc / cudaFree / allocating library call.  All calls on one device share the one array, which
 * keeps the previous call's contents (what must start at zero is cleared on job->stream at the start of every call):
 * two calls that overlap on the same GPU -- different streams, threads or processes' graphs replayed concurrently --
 * RACE; serialise them (one stream, or events).
 */
size_t pvv_static_workspace_bytes(void);
int pvv_v3_noalloc(const pvv_job *job, float confidence, int max_iter, float *out_kp);

/*
 * ransac_voting_layer_v5 (ransac_voting_gpu.py:769-864) = v3 + a per-keypoint confidence: the refined keypoints are
 * re-voted at the fixed threshold 0.999 (:854-856) and out_conf [b,vn] float32 = inlier count / tn (0 for skipped images).
 */
int pvv_v5(const pvv_job *job, float confidence, int max_iter, float *out_kp, float *out_conf, const pvv_debug *dbg);

/*
 * generate_hypothesis wrapper (ransac_voting_gpu.py:989-1040): the hypotheses and their inlier counts in the
 * reference's layouts, out_hyp [b,hn_total,vn,2] float32 and out_counts [b,hn_total,vn] int32 (device); zeros for
 * skipped images.  job->fg_rule = PVV_FG_BYTE.
 */
int pvv_hypotheses(const pvv_job *job, float *out_hyp, int32_t *out_counts, const pvv_debug *dbg);

/*
 * ransac_motion_voting (ransac_voting_gpu.py:966-987): out_kp [b,vn,2] float32 = mean over the foreground pixels of
 * (vertex + pixel coordinate); zeros for images without foreground.  Uses job->mask / vertex / strides / b,h,w,vn only
 * (fg rule PVV_FG_BYTE, no minimum, no subsampling); job->hn_total must be >= 1 and is otherwise ignored.
 */
int pvv_motion(const pvv_job *job, float *out_kp);

/*
 * estimate_voting_distribution_with_mean (ransac_voting_gpu.py:333-406):
 * mean [b,vn,2] float32 in, out_cov [b,vn,2,2] float32 out (device). job->hn_total = ceil(min_hyp/round_hyp)*round_hyp,
 * job->fg_rule must be PVV_FG_EQ1.
 */
int pvv_cov(const pvv_job *job, const float *mean, float *out_cov, const pvv_debug *dbg);

/*
 * Host-buffer variants (the reference-facing call with HOST memory): mask / vertex / idxs / selection / mean are host
 * pointers, outputs are host pointers; copies are issued on job->stream and the call returns after the result has
 * landed (it synchronises the stream).  `workspace` is still device memory, sized by pvv_workspace_bytes_host().
 * Pinned (cudaHostAlloc / cudaHostRegister) inputs are read in place over PCIe: only one byte per mask pixel and the
 * foreground of the vertex tensor cross the bus (contiguous masks are prepared by host worker threads -- see
 * pvv_set_host_workers -- and the foreground row spans of a [b,h,w,vn,2] view of an NCHW tensor, or of a contiguous
 * [b,h,w,vn,2] tensor, are staged in the workspace by a pull kernel; any other layout, and sparse foreground, is gathered
 * pixel by pixel by the compaction kernel).  Pageable inputs are staged with a full copy.  The workspace of a pinned call
 * holds a device image of the vertex tensor (same strides, only the foreground spans are written).
 * h2d_bytes / d2h_bytes (optional) receive the bytes that crossed the bus.
 */
size_t pvv_workspace_bytes_host(const pvv_job *job);
/*
 * Host worker threads of the host-buffer entry points.  They walk a contiguous mask piece by piece: a wide (int64 / int32 /
 * float32) mask is narrowed to one byte per pixel before it crosses PCIe (AVX-512 where the CPU has it; 8x fewer bytes for
 * the int64 mask torch.argmax produces), and the foreground extent of the rows is measured, which decides between the span
 * pull and the per-pixel gather.  n < 0 (default): three quarters of the hardware threads, at most 16; n = 0: no host preparation, the
 * mask is copied as it is and the foreground is gathered pixel by pixel; n > 0: that many.  The workers are parked threads
 * of the calling thread (created on its first call).  Measured on the B200 boxes (16 / 32 hardware threads): worth it from
 * 3 workers per rank upwards; a launcher that runs several ranks per box should divide the hardware threads between them
 * (the Python mirror does that from LOCAL_WORLD_SIZE).  Process-wide setting, may be changed at any time; returns the
 * previous value.
 */
int pvv_set_host_workers(int n);
int pvv_v3_host(const pvv_job *job, float confidence, int max_iter, float *out_kp_host,
                int64_t *h2d_bytes, int64_t *d2h_bytes);
int pvv_cov_host(const pvv_job *job, const float *mean_host, float *out_cov_host,
                 int64_t *h2d_bytes, int64_t *d2h_bytes);

/*
 * Primitive-compatible entry points with the reference's AoS layouts, so that the other wrappers of
 * ransac_voting_gpu.py (v1, v2, v4-v6, generate_hypothesis at :989) keep working unchanged:
 *   generate_hypothesis   (src/ransac_voting.cpp:20-31 -> ransac_voting_kernel.cu:11-86)
 *   voting_for_hypothesis (src/ransac_voting.cpp:41-55 -> ransac_voting_kernel.cu:88-167)
 * direct [tn,vn,2] f32, coords [tn,2] f32 (x,y), idxs [hn,vn,2] i32, hyp [hn,vn,2] f32 (written, zero for
 * degenerate pairs), inliers [hn,vn,tn] u8 (in place: only 1s are written, caller pre-zeroes).
 */
int pvv_generate_hypothesis(const float *direct, const float *coords, const int32_t *idxs, int tn, int vn, int hn,
                            float *hyp, void *stream);
int pvv_voting_for_hypothesis(const float *direct, const float *coords, const float *hyp, uint8_t *inliers, int tn,
                              int vn, int hn, float inlier_thresh, void *stream);
/* fused form of voting_for_hypothesis + torch.sum(inliers, 2): counts int32 [hn,vn], never materialises the matrix */
int pvv_vote_counts(const float *direct, const float *coords, const float *hyp, int tn, int vn, int hn,
                    float inlier_thresh, int32_t *counts, void *stream);

/*
 * The vanishing-point twins of the two primitives (src/ransac_voting.cpp:64-98 -> ransac_voting_kernel.cu:170-351),
 * so that the replacement module exports all four functions of the reference's pybind module (:102-107).
 * hyp is [hn,vn,3] f32: homogeneous intersection (x,y,z) of the two pixel lines, sign-flipped / zeroed as the
 * reference does (:219-223); the vote is |cos| > thresh between the direction and hyp.xy - coords*hyp.z with both
 * component products non-negative (:285-300).
 */
int pvv_generate_hypothesis_vanishing_point(const float *direct, const float *coords, const int32_t *idxs, int tn,
                                            int vn, int hn, float *hyp, void *stream);
int pvv_voting_for_hypothesis_vanishing_point(const float *direct, const float *coords, const float *hyp,
                                              uint8_t *inliers, int tn, int vn, int hn, float inlier_thresh,
                                              void *stream);

/*
 * Batched uncertainty-driven PnP -- the consumer of pvv_cov.  Replaces, per image,
 *   Evaluator.evaluate_uncertainty   lib/utils/evaluation_utils.py:164-183      (covariance -> 2x2 weights)
 *   uncertainty_pnp                  lib/utils/extend_utils/extend_utils.py:63-114 (start pose + cffi call)
 *   the Ceres problem                lib/utils/extend_utils/src/uncertainty_pnp.cpp:61-92
 * i.e. pose = argmin over (R,t) of sum_i || W_i (proj(K, R X_i + t) - x_i) ||^2.  All pointers are device pointers.
 *   pts2d           [b,vn,2]    keypoints (pvv_v3 / pvv_v5 output)
 *   weights_or_cov  mode 0: covariances [b,vn,2,2] (pvv_cov output): W_i = inv(sqrtm(cov_i)), zero weight when
 *                           cov_i[0,0] < 1e-6 or cov_i has a NaN (evaluation_utils.py:171-173)
 *                   mode 1: weights [b,vn,3] = wxx,wxy,wyy, what uncertainty_pnp() itself takes
 *   f64_inputs      0: pts2d / weights_or_cov are float32, 1: float64
 *   pts3d           [vn,3] float64 (+ pts3d_image_stride elements per image; 0 = one object for the whole batch)
 *   camera          [3,3] float64 row-major intrinsics (+ camera_image_stride; 0 = shared)
 *   out_pose        [b,3,4] float64 = [R|t], as uncertainty_pnp returns it
 *   out_info        optional [b,3] float64: final cost, LM iterations, status (0 converged, 1 iteration cap, 2 no start)
 */
int pvv_uncertainty_pnp(const void *pts2d, const void *weights_or_cov, int mode, int f64_inputs, const double *pts3d,
                        int64_t pts3d_image_stride, const double *camera, int64_t camera_image_stride, int b, int vn,
                        double *out_pose, double *out_info, void *stream);
/*
 * The same with a stop rule.  pvv_uncertainty_pnp runs the Levenberg-Marquardt iteration to convergence (the minimiser
 * Ceres approaches).  The reference sets no solver options (uncertainty_pnp.cpp:85-90), so Ceres stops at its defaults: 50
 * iterations (unsuccessful steps included) or an accepted step that lowers the cost by less than 1e-6 x cost.  Passing
 * max_iter = 50, function_tolerance = 1e-6 applies those two tests to this solver's own iteration; <= 0 switches a test
 * off.  The damping schedule is this library's, not Ceres' trust-region policy (Ceres itself cannot be built or linked in
 * this environment), so the early-stopped pose is close to, not identical with, what Ceres returns.
 */
int pvv_uncertainty_pnp_stop(const void *pts2d, const void *weights_or_cov, int mode, int f64_inputs, const double *pts3d,
                             int64_t pts3d_image_stride, const double *camera, int64_t camera_image_stride, int b, int vn,
                             int max_iter, double function_tolerance, double *out_pose, double *out_info, void *stream);

/* The library's draw streams, materialised (for feeding the oracle / the reference wrapper the same draws). */
int pvv_make_idxs(uint64_t seed, int64_t image_id_base, int b, int hn_total, int vn, const int32_t *tn /*device [b]*/,
                  int32_t *idxs /*device [b,hn_total,vn,2]*/, void *stream);
int pvv_make_selection(uint64_t seed, int64_t image_id_base, int b, int h, int w, float *selection /*device [b,h,w]*/,
                       void *stream);

/* FP32 roofline probe: runs an FFMA-only kernel for `iters` and returns achieved TFLOP/s (0 on error). */
double pvv_fp32_peak_tflops(int iters, void *stream);
/* diagnostic: kind 0 = scalar FFMA chains (same as above), kind 1 = packed fma.rn.f32x2 chains */
double pvv_probe_tflops(int kind, int iters, void *stream);

/* Number of kernel launches the last pvv_v3 / pvv_cov call on this thread enqueued. */
int pvv_last_launch_count(void);
int pvv_last_cuda_error(void);
const char *pvv_version(void);

#ifdef __cplusplus
}
#endif
#endif /* PVVOTE_H_ */
