/*
 * pvv_oracle.c -- CPU restatement of the reference RANSAC-voting kernels.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under pvnet_b200/ may import, link or
 * call this file; only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs use it, and only as the checker or the
 * reported CPU baseline.
 *
 * It follows, function by function (paths relative to /root/reference):
 *   orc_generate_hypothesis   <- lib/ransac_voting_gpu_layer/src/ransac_voting_kernel.cu:11-49
 *                                (+ at::zeros at :75: degenerate pairs stay 0,0)
 *   orc_voting_for_hypothesis <- lib/ransac_voting_gpu_layer/src/ransac_voting_kernel.cu:88-126
 *   orc_vote_counts           <- same kernel + torch.sum(inlier, 2)
 *                                (lib/ransac_voting_gpu_layer/ransac_voting_gpu.py:563-567)
 *
 * Floating point: the reference is built with nvcc's defaults (-fmad=true,
 * IEEE sqrt/div).  The contraction pattern below is the one nvcc 12.9 emits
 * for sm_100a (cuobjdump -sass of the reference .cu, see DESIGN.md "Reference
 * arithmetic"); every fused multiply-add is an explicit fmaf() and this file
 * must be compiled with -ffp-contract=off so nothing else fuses.  Both 1e-6
 * literals in the reference are double: the float operand is promoted before
 * the compare.
 *
 * Parity pin: the reference ships no golden vectors for this path
 * (SURVEY.md section 4).  The pin is (a) tests/golden/ fixtures produced by
 * running the reference's own Python wrapper over this file's kernels
 * (tests/golden/make_golden.py) and (b) on the GPU box, a bitwise A/B against
 * the reference's own CUDA kernels built into oracle/_ref/.
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

int orc_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* torchrun exports OMP_NUM_THREADS=1 to its workers; the timed CPU baseline asks for the host's cores explicitly */
void orc_set_num_threads(int n)
{
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/* direct [tn,vn,2], coords [tn,2] (x,y), idxs [hn,vn,2] -> hyp [hn,vn,2] */
void orc_generate_hypothesis(const float *direct, const float *coords,
                             const int32_t *idxs, int tn, int vn, int hn,
                             float *hyp)
{
    (void)tn;
    memset(hyp, 0, sizeof(float) * (size_t)hn * vn * 2);
    for (int hi = 0; hi < hn; ++hi) {
        for (int vi = 0; vi < vn; ++vi) {
            const int t0 = idxs[(hi * vn + vi) * 2];
            const int t1 = idxs[(hi * vn + vi) * 2 + 1];
            const float d0x = direct[((size_t)t0 * vn + vi) * 2];
            const float d0y = direct[((size_t)t0 * vn + vi) * 2 + 1];
            const float d1x = direct[((size_t)t1 * vn + vi) * 2];
            const float d1y = direct[((size_t)t1 * vn + vi) * 2 + 1];
            const float cx0 = coords[(size_t)t0 * 2], cy0 = coords[(size_t)t0 * 2 + 1];
            const float cx1 = coords[(size_t)t1 * 2], cy1 = coords[(size_t)t1 * 2 + 1];

            /* line normals are (dy,-dx); det_y = nx1*ny0 - nx0*ny1 */
            const float p = d0y * d1x;
            const float q = d0x * d1y;
            const float det_y = p - q;
            if (fabs((double)det_y) < 1e-6) continue;
            const float det_x = q - p;
            if (fabs((double)det_x) < 1e-6) continue;

            const float s1 = fmaf(d1y, cx1, -(d1x * cy1));
            const float s0 = fmaf(d0y, cx0, -(d0x * cy0));
            const float num_y = fmaf(d1y, s0, -(d0y * s1));
            const float num_x = fmaf(d0x, s1, -(d1x * s0));
            hyp[(hi * vn + vi) * 2] = num_x / det_x;
            hyp[(hi * vn + vi) * 2 + 1] = num_y / det_y;
        }
    }
}

/* one (hypothesis, keypoint, pixel) test, reference arithmetic */
static inline int orc_test(float nx, float ny, float cx, float cy, float hx,
                           float hy, float thresh)
{
    const float dx = hx - cx;
    const float dy = hy - cy;
    const float n1 = sqrtf(fmaf(nx, nx, ny * ny));
    const float n2 = sqrtf(fmaf(dx, dx, dy * dy));
    if ((double)n1 < 1e-6 || (double)n2 < 1e-6) return 0;
    const float den = n2 * n1;
    const float num = fmaf(nx, dx, ny * dy);
    return (num / den) > thresh;
}

/* scalar entry so tests can probe single cases */
int orc_inlier_test(float nx, float ny, float cx, float cy, float hx, float hy,
                    float thresh)
{
    return orc_test(nx, ny, cx, cy, hx, hy, thresh);
}

/* In-place like the reference: only writes 1s, caller pre-zeroes inliers. */
void orc_voting_for_hypothesis(const float *direct, const float *coords,
                               const float *hyp, uint8_t *inliers, int tn,
                               int vn, int hn, float thresh)
{
#pragma omp parallel for schedule(static)
    for (int hi = 0; hi < hn; ++hi) {
        for (int vi = 0; vi < vn; ++vi) {
            const float hx = hyp[(hi * vn + vi) * 2];
            const float hy = hyp[(hi * vn + vi) * 2 + 1];
            uint8_t *row = inliers + ((size_t)hi * vn + vi) * tn;
            for (int ti = 0; ti < tn; ++ti) {
                if (orc_test(direct[((size_t)ti * vn + vi) * 2],
                             direct[((size_t)ti * vn + vi) * 2 + 1],
                             coords[(size_t)ti * 2], coords[(size_t)ti * 2 + 1],
                             hx, hy, thresh))
                    row[ti] = 1;
            }
        }
    }
}

/* counts[hi,vi] = sum_ti inlier[hi,vi,ti] without materialising the matrix */
void orc_vote_counts(const float *direct, const float *coords,
                     const float *hyp, int tn, int vn, int hn, float thresh,
                     int64_t *counts)
{
#pragma omp parallel for schedule(static) collapse(2)
    for (int hi = 0; hi < hn; ++hi) {
        for (int vi = 0; vi < vn; ++vi) {
            const float hx = hyp[(hi * vn + vi) * 2];
            const float hy = hyp[(hi * vn + vi) * 2 + 1];
            int64_t c = 0;
            for (int ti = 0; ti < tn; ++ti)
                c += orc_test(direct[((size_t)ti * vn + vi) * 2],
                              direct[((size_t)ti * vn + vi) * 2 + 1],
                              coords[(size_t)ti * 2], coords[(size_t)ti * 2 + 1],
                              hx, hy, thresh);
            counts[hi * vn + vi] = c;
        }
    }
}

/* ---- vanishing-point twins: lib/ransac_voting_gpu_layer/src/ransac_voting_kernel.cu:170-227 and :258-301 ----
 * Rounding sequence as compiled by the reference build (nvcc defaults, fma contraction on; read off the sm_100a
 * SASS of oracle/_ref/ransac_voting_kernel.o): products that feed a subtraction are rounded, the subtraction is the
 * fma; num of the cosine is the plain sum of the two rounded products that are also val_x / val_y. */
void orc_generate_hypothesis_vanishing_point(const float *direct, const float *coords, const int32_t *idxs, int tn,
                                             int vn, int hn, float *hyp /* [hn,vn,3] */)
{
    (void)tn;
    for (int hi = 0; hi < hn; ++hi) {
        for (int vi = 0; vi < vn; ++vi) {
            const int t0 = idxs[(hi * vn + vi) * 2], t1 = idxs[(hi * vn + vi) * 2 + 1];
            const float dx0 = direct[((size_t)t0 * vn + vi) * 2], dy0 = direct[((size_t)t0 * vn + vi) * 2 + 1];
            const float dx1 = direct[((size_t)t1 * vn + vi) * 2], dy1 = direct[((size_t)t1 * vn + vi) * 2 + 1];
            const float cx0 = coords[(size_t)t0 * 2], cy0 = coords[(size_t)t0 * 2 + 1];
            const float cx1 = coords[(size_t)t1 * 2], cy1 = coords[(size_t)t1 * 2 + 1];
            const float lz1 = fmaf(dx1, cy1, -(dy1 * cx1));
            const float lz0 = fmaf(dx0, cy0, -(dy0 * cx0));
            float z = fmaf(dx0, dy1, -(dy0 * dx1));
            float x = fmaf(dx1, lz0, -(dx0 * lz1));
            float y = fmaf(dy1, lz0, -(dy0 * lz1));
            const float vx0 = dx0 * fmaf(-cx0, z, x), vx1 = dx1 * fmaf(-cx1, z, x);
            const float vy0 = dy0 * fmaf(-cy0, z, y), vy1 = dy1 * fmaf(-cy1, z, y);
            if (vx0 < 0 && vx1 < 0 && vy0 < 0 && vy1 < 0) { x = -x; y = -y; z = -z; }
            if (vx0 * vx1 < 0 || vy0 * vy1 < 0) { x = 0.f; y = 0.f; z = 0.f; }
            float *o = hyp + ((size_t)hi * vn + vi) * 3;
            o[0] = x; o[1] = y; o[2] = z;
        }
    }
}

void orc_voting_for_hypothesis_vanishing_point(const float *direct, const float *coords, const float *hyp,
                                               uint8_t *inliers, int tn, int vn, int hn, float thresh)
{
#pragma omp parallel for schedule(static)
    for (int hi = 0; hi < hn; ++hi) {
        for (int vi = 0; vi < vn; ++vi) {
            const float *h = hyp + ((size_t)hi * vn + vi) * 3;
            uint8_t *row = inliers + ((size_t)hi * vn + vi) * tn;
            for (int ti = 0; ti < tn; ++ti) {
                const float dx = direct[((size_t)ti * vn + vi) * 2], dy = direct[((size_t)ti * vn + vi) * 2 + 1];
                const float fx = fmaf(-coords[(size_t)ti * 2], h[2], h[0]);
                const float fy = fmaf(-coords[(size_t)ti * 2 + 1], h[2], h[1]);
                const float n1 = sqrtf(fmaf(dx, dx, dy * dy));
                const float n2 = sqrtf(fmaf(fx, fx, fy * fy));
                if ((double)n1 < 1e-6 || (double)n2 < 1e-6) continue;
                const float vx = dx * fx, vy = dy * fy;
                if (vx < 0 || vy < 0) continue;
                if (fabsf((vx + vy) / (n2 * n1)) > thresh) row[ti] = 1;
            }
        }
    }
}
