// pvv_pnp.cu -- N2: batched uncertainty-driven PnP, the consumer of the covariance path.
//
// Replaces, for a whole batch in one launch and without the .cpu() hop,
//   Evaluator.evaluate_uncertainty      lib/utils/evaluation_utils.py:164-183   (covariance -> 2x2 weights)
//   uncertainty_pnp                     lib/utils/extend_utils/extend_utils.py:63-114
//   the Ceres problem it solves         lib/utils/extend_utils/src/uncertainty_pnp.cpp:7-92
// i.e. per image:   minimise over (R, t)   sum_i || W_i (proj(K, R X_i + t) - x_i) ||^2,   W_i = inv(sqrtm(cov_i)).
//
// One warp per image, lane = keypoint; everything in float64 (Ceres and the reference wrapper work in double).
//   weights   closed-form inverse square root of the symmetric 2x2 covariance (what scipy.linalg.sqrtm + inv give for an
//             SPD matrix); a keypoint with cov[0,0] < 1e-6 or any NaN gets zero weight (evaluation_utils.py:171-173)
//   start     scaled-orthographic pose from all keypoints (one 3x3 solve) -- the reference starts from OpenCV's P3P on
//             the four heaviest keypoints (extend_utils.py:83-87); either is only a starting point for the refinement
//   refine    Levenberg-Marquardt on SE(3) (left-multiplicative rotation update), warp-shuffle reductions of the 6x6
//             normal equations, Cholesky solve in registers, run to convergence (the reference: Ceres defaults,
//             DENSE_SCHUR, 50 iterations, function tolerance 1e-6)
// Parity is asserted on the converged pose (tests/test_gpu_parity.py::test_uncertainty_pnp_*); DESIGN.md lists the
// deviations (start pose, stopping rule, the pn == 4 shortcut).
#include <math.h>

#include "pvv_internal.cuh"

namespace {

constexpr int kPnpMaxIter = 100;

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// inverse of the principal square root of a symmetric 2x2 matrix [[a,b],[b,c]]; returns false if it is not SPD
__device__ __forceinline__ bool inv_sqrt_sym2(double a, double b, double c, double &wxx, double &wxy, double &wyy)
{
    const double det = a * c - b * b;
    if (!(det > 0.0) || !(a > 0.0)) return false;
    const double s = sqrt(det), t = sqrt(a + c + 2.0 * s);
    // sqrt(A) = (A + s I) / t ;  inverse of a symmetric 2x2
    const double ra = (a + s) / t, rb = b / t, rc = (c + s) / t;
    const double rdet = ra * rc - rb * rb;  // = s
    if (!(rdet > 0.0) || !isfinite(rdet)) return false;
    wxx = rc / rdet; wxy = -rb / rdet; wyy = ra / rdet;
    return true;
}

// solve the SPD 6x6 system A x = g in place (Cholesky); returns false when A is not positive definite
__device__ __forceinline__ bool chol6(double (&A)[6][6], double (&g)[6])
{
#pragma unroll
    for (int j = 0; j < 6; ++j) {
        double d = A[j][j];
#pragma unroll
        for (int k = 0; k < j; ++k) d -= A[j][k] * A[j][k];
        if (!(d > 0.0)) return false;
        d = sqrt(d);
        A[j][j] = d;
#pragma unroll
        for (int i = j + 1; i < 6; ++i) {
            double v = A[i][j];
#pragma unroll
            for (int k = 0; k < j; ++k) v -= A[i][k] * A[j][k];
            A[i][j] = v / d;
        }
    }
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        double v = g[i];
#pragma unroll
        for (int k = 0; k < i; ++k) v -= A[i][k] * g[k];
        g[i] = v / A[i][i];
    }
#pragma unroll
    for (int i = 5; i >= 0; --i) {
        double v = g[i];
#pragma unroll
        for (int k = i + 1; k < 6; ++k) v -= A[k][i] * g[k];
        g[i] = v / A[i][i];
    }
    return true;
}

// R <- exp([w]x) R  (Rodrigues), re-orthonormalised by construction of exp
__device__ __forceinline__ void rot_update(const double (&w)[3], const double (&R)[9], double (&Rn)[9])
{
    const double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2];
    double a, b;  // exp = I + a [w]x + b [w]x^2
    if (th2 > 1e-16) {
        const double th = sqrt(th2);
        a = sin(th) / th;
        b = (1.0 - cos(th)) / th2;
    } else {
        a = 1.0 - th2 / 6.0;
        b = 0.5 - th2 / 24.0;
    }
    double E[9];
    E[0] = 1.0 - b * (w[1] * w[1] + w[2] * w[2]); E[1] = -a * w[2] + b * w[0] * w[1]; E[2] = a * w[1] + b * w[0] * w[2];
    E[3] = a * w[2] + b * w[0] * w[1]; E[4] = 1.0 - b * (w[0] * w[0] + w[2] * w[2]); E[5] = -a * w[0] + b * w[1] * w[2];
    E[6] = -a * w[1] + b * w[0] * w[2]; E[7] = a * w[0] + b * w[1] * w[2]; E[8] = 1.0 - b * (w[0] * w[0] + w[1] * w[1]);
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) Rn[i * 3 + j] = E[i * 3] * R[j] + E[i * 3 + 1] * R[3 + j] + E[i * 3 + 2] * R[6 + j];
}

struct pnp_point { double X[3], x, y, wxx, wxy, wyy; };

// weighted reprojection residual of one keypoint and (optionally) its 2x6 Jacobian w.r.t. (delta rotation, delta t)
__device__ __forceinline__ void residual(const pnp_point &p, const double (&R)[9], const double (&t)[3], double fx, double fy,
                                         double cx, double cy, double &r0, double &r1, double (*J)[6])
{
    const double Y0 = R[0] * p.X[0] + R[1] * p.X[1] + R[2] * p.X[2];
    const double Y1 = R[3] * p.X[0] + R[4] * p.X[1] + R[5] * p.X[2];
    const double Y2 = R[6] * p.X[0] + R[7] * p.X[1] + R[8] * p.X[2];
    const double P0 = Y0 + t[0], P1 = Y1 + t[1], P2 = Y2 + t[2];
    const double iz = 1.0 / P2;
    const double dx = fx * P0 * iz + cx - p.x, dy = fy * P1 * iz + cy - p.y;
    r0 = p.wxx * dx + p.wxy * dy;
    r1 = p.wxy * dx + p.wyy * dy;
    if (!J) return;
    // d proj / d P
    const double a00 = fx * iz, a02 = -fx * P0 * iz * iz, a11 = fy * iz, a12 = -fy * P1 * iz * iz;
    // d P / d (delta rotation) = -[Y]x  (left update: P = exp(w) Y + t), d P / d t = I
    double D[2][6];
    D[0][0] = a02 * Y1;              D[0][1] = a00 * Y2 - a02 * Y0;   D[0][2] = -a00 * Y1;
    D[1][0] = -a11 * Y2 + a12 * Y1;  D[1][1] = -a12 * Y0;             D[1][2] = a11 * Y0;
    D[0][3] = a00; D[0][4] = 0.0; D[0][5] = a02;
    D[1][3] = 0.0; D[1][4] = a11; D[1][5] = a12;
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        J[0][k] = p.wxx * D[0][k] + p.wxy * D[1][k];
        J[1][k] = p.wxy * D[0][k] + p.wyy * D[1][k];
    }
}

// mode 0: `w_or_cov` holds covariances [b,vn,2,2] (evaluate_uncertainty); mode 1: weights [b,vn,3] = wxx,wxy,wyy (uncertainty_pnp)
// T: element type of pts2d and w_or_cov (float32 straight from the voting layer, or float64 like the reference's numpy path)
template <typename T>
__global__ void __launch_bounds__(128) k_uncertainty_pnp(const T *__restrict__ pts2d, const T *__restrict__ w_or_cov,
                                                         int mode, const double *__restrict__ pts3d, long long pts3d_stride,
                                                         const double *__restrict__ Kmat, long long k_stride, int b, int vn,
                                                         double *__restrict__ out_pose, double *__restrict__ out_info,
                                                         int max_iter, double function_tolerance)
{
    const int img = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (img >= b) return;
    const double *Km = Kmat + img * k_stride;
    const double fx = Km[0], fy = Km[4], cx = Km[2], cy = Km[5];

    // ---- this lane's keypoint(s): vn <= 32 in PVNet (9); more are folded in strides of 32 ----
    constexpr int MAXP = PVV_MAX_VN / 32;  // 2
    pnp_point pt[MAXP];
    int np = 0;
    for (int vi = lane; vi < vn && np < MAXP; vi += 32, ++np) {
        pnp_point &p = pt[np];
        const double *X = pts3d + img * pts3d_stride + vi * 3;
        p.X[0] = X[0]; p.X[1] = X[1]; p.X[2] = X[2];
        const long long o = (long long)img * vn + vi;
        p.x = pts2d[o * 2]; p.y = pts2d[o * 2 + 1];
        p.wxx = p.wxy = p.wyy = 0.0;
        if (mode == 0) {
            const T *c = w_or_cov + o * 4;
            const bool bad = c[0] < (T)1e-6 || c[0] != c[0] || c[1] != c[1] || c[2] != c[2] || c[3] != c[3];
            // sqrtm of the (numerically slightly asymmetric) matrix: symmetrise the off-diagonal
            if (!bad) inv_sqrt_sym2((double)c[0], 0.5 * ((double)c[1] + (double)c[2]), (double)c[3], p.wxx, p.wxy, p.wyy);
        } else {
            const T *w = w_or_cov + o * 3;
            p.wxx = w[0]; p.wxy = w[1]; p.wyy = w[2];
        }
        if (!isfinite(p.x) || !isfinite(p.y)) { p.wxx = p.wxy = p.wyy = 0.0; p.x = p.y = 0.0; }
    }

    // ---- start pose: scaled orthographic projection, least squares over the weighted-in keypoints ----
    double n = 0, mX[3] = {0, 0, 0}, mu = 0, mv = 0;
    for (int k = 0; k < np; ++k) {
        const bool use = pt[k].wxx != 0.0 || pt[k].wyy != 0.0;
        if (use) { n += 1; mX[0] += pt[k].X[0]; mX[1] += pt[k].X[1]; mX[2] += pt[k].X[2];
                   mu += (pt[k].x - cx) / fx; mv += (pt[k].y - cy) / fy; }
    }
    n = warp_sum(n);
#pragma unroll
    for (int i = 0; i < 3; ++i) mX[i] = warp_sum(mX[i]);
    mu = warp_sum(mu); mv = warp_sum(mv);
    double R[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, t[3] = {0, 0, 1};
    bool ok = n >= 3.5;  // at least 4 usable keypoints
    if (ok) {
        const double inv_n = 1.0 / n;
#pragma unroll
        for (int i = 0; i < 3; ++i) mX[i] *= inv_n;
        mu *= inv_n; mv *= inv_n;
        double M[6] = {0, 0, 0, 0, 0, 0}, bu[3] = {0, 0, 0}, bv[3] = {0, 0, 0};  // M: xx xy xz yy yz zz
        for (int k = 0; k < np; ++k) {
            if (!(pt[k].wxx != 0.0 || pt[k].wyy != 0.0)) continue;
            const double d0 = pt[k].X[0] - mX[0], d1 = pt[k].X[1] - mX[1], d2 = pt[k].X[2] - mX[2];
            const double du = (pt[k].x - cx) / fx - mu, dv = (pt[k].y - cy) / fy - mv;
            M[0] += d0 * d0; M[1] += d0 * d1; M[2] += d0 * d2; M[3] += d1 * d1; M[4] += d1 * d2; M[5] += d2 * d2;
            bu[0] += d0 * du; bu[1] += d1 * du; bu[2] += d2 * du;
            bv[0] += d0 * dv; bv[1] += d1 * dv; bv[2] += d2 * dv;
        }
#pragma unroll
        for (int i = 0; i < 6; ++i) M[i] = warp_sum(M[i]);
#pragma unroll
        for (int i = 0; i < 3; ++i) { bu[i] = warp_sum(bu[i]); bv[i] = warp_sum(bv[i]); }
        // ridge: coplanar keypoints leave M singular; the refinement still has a usable start
        const double ridge = 1e-9 * (M[0] + M[3] + M[5]) + 1e-300;
        M[0] += ridge; M[3] += ridge; M[5] += ridge;
        // inverse of the symmetric 3x3 via cofactors
        const double c00 = M[3] * M[5] - M[4] * M[4], c01 = M[2] * M[4] - M[1] * M[5], c02 = M[1] * M[4] - M[2] * M[3];
        const double c11 = M[0] * M[5] - M[2] * M[2], c12 = M[1] * M[2] - M[0] * M[4], c22 = M[0] * M[3] - M[1] * M[1];
        const double det = M[0] * c00 + M[1] * c01 + M[2] * c02;
        double a[3], bb[3];
        a[0] = (c00 * bu[0] + c01 * bu[1] + c02 * bu[2]) / det;
        a[1] = (c01 * bu[0] + c11 * bu[1] + c12 * bu[2]) / det;
        a[2] = (c02 * bu[0] + c12 * bu[1] + c22 * bu[2]) / det;
        bb[0] = (c00 * bv[0] + c01 * bv[1] + c02 * bv[2]) / det;
        bb[1] = (c01 * bv[0] + c11 * bv[1] + c12 * bv[2]) / det;
        bb[2] = (c02 * bv[0] + c12 * bv[1] + c22 * bv[2]) / det;
        const double na = sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]), nb = sqrt(bb[0] * bb[0] + bb[1] * bb[1] + bb[2] * bb[2]);
        ok = isfinite(na) && isfinite(nb) && na > 0.0 && nb > 0.0;
        if (ok) {
            const double s = sqrt(na * nb);
            double r1[3] = {a[0] / na, a[1] / na, a[2] / na};
            const double dot = (bb[0] * r1[0] + bb[1] * r1[1] + bb[2] * r1[2]);
            double r2[3] = {bb[0] - dot * r1[0], bb[1] - dot * r1[1], bb[2] - dot * r1[2]};
            const double n2 = sqrt(r2[0] * r2[0] + r2[1] * r2[1] + r2[2] * r2[2]);
            ok = n2 > 1e-12 * nb;
            if (ok) {
                r2[0] /= n2; r2[1] /= n2; r2[2] /= n2;
                R[0] = r1[0]; R[1] = r1[1]; R[2] = r1[2];
                R[3] = r2[0]; R[4] = r2[1]; R[5] = r2[2];
                R[6] = r1[1] * r2[2] - r1[2] * r2[1]; R[7] = r1[2] * r2[0] - r1[0] * r2[2]; R[8] = r1[0] * r2[1] - r1[1] * r2[0];
                const double tz = 1.0 / s;
                t[0] = mu * tz - (R[0] * mX[0] + R[1] * mX[1] + R[2] * mX[2]);
                t[1] = mv * tz - (R[3] * mX[0] + R[4] * mX[1] + R[5] * mX[2]);
                t[2] = tz - (R[6] * mX[0] + R[7] * mX[1] + R[8] * mX[2]);
            }
        }
    }

    // ---- Levenberg-Marquardt ----
    auto cost_of = [&](const double (&Rc)[9], const double (&tc)[3]) {
        double c = 0.0;
        for (int k = 0; k < np; ++k) {
            double r0, r1;
            residual(pt[k], Rc, tc, fx, fy, cx, cy, r0, r1, nullptr);
            c += r0 * r0 + r1 * r1;
        }
        return warp_sum(c);
    };
    double cost = cost_of(R, t), lambda = 1e-4;
    int it = 0, status = ok ? 0 : 2;  // 0 converged, 1 iteration cap, 2 no usable start (pose left at the fallback)
    // Stop rule.  Default (max_iter <= 0, function_tolerance <= 0): to convergence.  With the reference's solver options
    // (uncertainty_pnp.cpp:85-90 sets none, so Ceres' defaults hold: 50 iterations, function tolerance 1e-6) the loop ends
    // after max_iter damping trials or when an accepted step lowers the cost by less than function_tolerance x cost -- the
    // test Ceres applies -- which is what makes "the pose the reference returns" differ from the minimiser on flat problems.
    const int iter_cap = max_iter > 0 ? min(max_iter, kPnpMaxIter) : kPnpMaxIter;
    int trials = 0;
    for (; ok && it < iter_cap; ++it) {
        double H[21], g[6];
#pragma unroll
        for (int i = 0; i < 21; ++i) H[i] = 0.0;
#pragma unroll
        for (int i = 0; i < 6; ++i) g[i] = 0.0;
        for (int k = 0; k < np; ++k) {
            double r0, r1, J[2][6];
            residual(pt[k], R, t, fx, fy, cx, cy, r0, r1, J);
            int q = 0;
#pragma unroll
            for (int i = 0; i < 6; ++i) {
                g[i] += J[0][i] * r0 + J[1][i] * r1;
#pragma unroll
                for (int j = i; j < 6; ++j) H[q++] += J[0][i] * J[0][j] + J[1][i] * J[1][j];
            }
        }
#pragma unroll
        for (int i = 0; i < 21; ++i) H[i] = warp_sum(H[i]);
#pragma unroll
        for (int i = 0; i < 6; ++i) g[i] = warp_sum(g[i]);
        double gmax = 0.0;
#pragma unroll
        for (int i = 0; i < 6; ++i) gmax = fmax(gmax, fabs(g[i]));
        if (!(gmax > 1e-14 * (1.0 + cost))) break;  // stationary
        bool improved = false;
        for (int tries = 0; tries < 12 && !improved; ++tries) {
            if (max_iter > 0 && trials++ >= max_iter) { status = 1; goto done; }   // Ceres counts unsuccessful steps too
            double A[6][6], d[6];
            int q = 0;
#pragma unroll
            for (int i = 0; i < 6; ++i)
#pragma unroll
                for (int j = i; j < 6; ++j) { A[j][i] = H[q]; A[i][j] = H[q]; ++q; }
#pragma unroll
            for (int i = 0; i < 6; ++i) { A[i][i] += lambda * (A[i][i] + 1e-12); d[i] = -g[i]; }
            if (chol6(A, d)) {
                double Rn[9], tn[3] = {t[0] + d[3], t[1] + d[4], t[2] + d[5]};
                const double w[3] = {d[0], d[1], d[2]};
                rot_update(w, R, Rn);
                const double cn = cost_of(Rn, tn);
                if (cn < cost || (cn == cost && cn == 0.0)) {
                    const double rel = (cost - cn) / fmax(cost, 1e-300);
                    double step2 = 0.0;
#pragma unroll
                    for (int i = 0; i < 6; ++i) step2 += d[i] * d[i];
#pragma unroll
                    for (int i = 0; i < 9; ++i) R[i] = Rn[i];
                    t[0] = tn[0]; t[1] = tn[1]; t[2] = tn[2];
                    cost = cn;
                    lambda = fmax(lambda / 3.0, 1e-12);
                    improved = true;
                    if (rel < 1e-15 || step2 < 1e-26) { ++it; goto done; }
                    if (function_tolerance > 0.0 && rel < function_tolerance) { ++it; goto done; }
                    continue;
                }
            }
            lambda *= 4.0;
        }
        if (!improved) break;  // no downhill step at any damping: at the minimum to working precision
    }
done:
    if (status == 0 && it >= iter_cap) status = 1;
    if (lane == 0) {
        double *o = out_pose + (long long)img * 12;
        o[0] = R[0]; o[1] = R[1]; o[2] = R[2]; o[3] = t[0];
        o[4] = R[3]; o[5] = R[4]; o[6] = R[5]; o[7] = t[1];
        o[8] = R[6]; o[9] = R[7]; o[10] = R[8]; o[11] = t[2];
        if (out_info) { out_info[img * 3] = cost; out_info[img * 3 + 1] = (double)it; out_info[img * 3 + 2] = (double)status; }
    }
}

}  // namespace

extern "C" int pvv_uncertainty_pnp(const void *pts2d, const void *weights_or_cov, int mode, int f64_inputs,
                                   const double *pts3d, int64_t pts3d_image_stride, const double *camera,
                                   int64_t camera_image_stride, int b, int vn, double *out_pose, double *out_info,
                                   void *stream)
{
    return pvv_uncertainty_pnp_stop(pts2d, weights_or_cov, mode, f64_inputs, pts3d, pts3d_image_stride, camera,
                                    camera_image_stride, b, vn, 0, 0.0, out_pose, out_info, stream);
}

extern "C" int pvv_uncertainty_pnp_stop(const void *pts2d, const void *weights_or_cov, int mode, int f64_inputs,
                                        const double *pts3d, int64_t pts3d_image_stride, const double *camera,
                                        int64_t camera_image_stride, int b, int vn, int max_iter, double function_tolerance,
                                        double *out_pose, double *out_info, void *stream)
{
    if (!pts2d || !weights_or_cov || !pts3d || !camera || !out_pose || b < 0 || vn < 4 || (mode != 0 && mode != 1))
        return PVV_ERR_INVALID;
    if (vn > PVV_MAX_VN) return PVV_ERR_UNSUPPORTED;
    if (b == 0) return PVV_OK;
    const dim3 grid((b + 3) / 4);
    if (f64_inputs)
        k_uncertainty_pnp<double><<<grid, 128, 0, (cudaStream_t)stream>>>((const double *)pts2d, (const double *)weights_or_cov,
                                                                          mode, pts3d, pts3d_image_stride, camera,
                                                                          camera_image_stride, b, vn, out_pose, out_info,
                                                                          max_iter, function_tolerance);
    else
        k_uncertainty_pnp<float><<<grid, 128, 0, (cudaStream_t)stream>>>((const float *)pts2d, (const float *)weights_or_cov,
                                                                         mode, pts3d, pts3d_image_stride, camera,
                                                                         camera_image_stride, b, vn, out_pose, out_info,
                                                                         max_iter, function_tolerance);
    return cudaGetLastError() == cudaSuccess ? PVV_OK : PVV_ERR_CUDA;
}
