// pvv_primitives.cu -- primitive-compatible entry points with the reference's AoS tensor contracts
// (direct [tn,vn,2], coords [tn,2], idxs/hyp [hn,vn,2], inliers [hn,vn,tn] u8), so that the remaining wrappers
// of ransac_voting_gpu.py (v1, v2, v4-v6, generate_hypothesis :989) keep working on top of this library.
//   pvv_generate_hypothesis   <- ransac_voting.cpp:20-31 / ransac_voting_kernel.cu:11-86
//   pvv_voting_for_hypothesis <- ransac_voting.cpp:41-55 / ransac_voting_kernel.cu:88-167
//   pvv_vote_counts           <- the same + torch.sum(inliers, 2), fused
// Unlike the reference these run on the caller's stream and return a status instead of exit()ing.
#include "pvv_internal.cuh"

namespace {

__global__ void __launch_bounds__(256) k_prim_generate(const float *__restrict__ direct, const float *__restrict__ coords,
                                                       const int32_t *__restrict__ idxs, int tn, int vn, int hn,
                                                       float *__restrict__ hyp)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= hn * vn) return;
    const int vi = e % vn;
    const int t0 = idxs[e * 2], t1 = idxs[e * 2 + 1];
    float x = 0.f, y = 0.f;
    if ((unsigned)t0 < (unsigned)tn && (unsigned)t1 < (unsigned)tn) {
        const float2 d0 = *(const float2 *)(direct + ((long long)t0 * vn + vi) * 2);
        const float2 d1 = *(const float2 *)(direct + ((long long)t1 * vn + vi) * 2);
        const float2 c0 = *(const float2 *)(coords + (long long)t0 * 2);
        const float2 c1 = *(const float2 *)(coords + (long long)t1 * 2);
        float hx, hy;
        if (pvv_intersect(d0.x, d0.y, c0.x, c0.y, d1.x, d1.y, c1.x, c1.y, hx, hy)) { x = hx; y = hy; }
    }
    hyp[e * 2] = x;
    hyp[e * 2 + 1] = y;
}

// grid.y = hn*vn rows, grid.x covers tn; consecutive threads -> consecutive pixels (coalesced u8 stores)
__global__ void __launch_bounds__(256) k_prim_vote(const float *__restrict__ direct, const float *__restrict__ coords,
                                                   const float *__restrict__ hyp, uint8_t *__restrict__ inliers, int tn,
                                                   int vn, int hn, float thresh)
{
    const int rowi = blockIdx.y;  // hi*vn + vi
    const int vi = rowi % vn;
    const float hx = hyp[rowi * 2], hy = hyp[rowi * 2 + 1];
    for (int ti = blockIdx.x * blockDim.x + threadIdx.x; ti < tn; ti += gridDim.x * blockDim.x) {
        const float2 dir = *(const float2 *)(direct + ((long long)ti * vn + vi) * 2);
        const float2 c = *(const float2 *)(coords + (long long)ti * 2);
        if (pvv_exact_test(dir.x, dir.y, pvv_norm1(dir.x, dir.y), c.x, c.y, hx, hy, thresh))
            inliers[(long long)rowi * tn + ti] = 1;
    }
}

__global__ void __launch_bounds__(256) k_prim_counts(const float *__restrict__ direct, const float *__restrict__ coords,
                                                     const float *__restrict__ hyp, int tn, int vn, int hn, float thresh,
                                                     int32_t *__restrict__ counts)
{
    const int rowi = blockIdx.x;
    const int vi = rowi % vn;
    const float hx = hyp[rowi * 2], hy = hyp[rowi * 2 + 1];
    int c = 0;
    for (int ti = threadIdx.x; ti < tn; ti += blockDim.x) {
        const float2 dir = *(const float2 *)(direct + ((long long)ti * vn + vi) * 2);
        const float2 p = *(const float2 *)(coords + (long long)ti * 2);
        c += pvv_exact_test(dir.x, dir.y, pvv_norm1(dir.x, dir.y), p.x, p.y, hx, hy, thresh);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    __shared__ int s[8];
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int i = 0; i < 8; ++i) t += s[i];
        counts[rowi] = t;
    }
}

// ---- vanishing-point twins (ransac_voting_kernel.cu:170-351): hypotheses are homogeneous points (x, y, z) ----
// Rounding sequence of the reference build (nvcc 12.9 defaults, sm_100a SASS of oracle/_ref/ransac_voting_kernel.o):
//   lz_k = fma(dx_k, cy_k, -rn(dy_k*cx_k));  z = fma(dx0, dy1, -rn(dy0*dx1));  x = fma(dx1, lz0, -rn(dx0*lz1));
//   y = fma(dy1, lz0, -rn(dy0*lz1));  val_x_k = rn(dx_k * fma(-cx_k, z, x));  val_y_k = rn(dy_k * fma(-cy_k, z, y))
__global__ void __launch_bounds__(256) k_prim_generate_vp(const float *__restrict__ direct, const float *__restrict__ coords,
                                                          const int32_t *__restrict__ idxs, int tn, int vn, int hn,
                                                          float *__restrict__ hyp)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= hn * vn) return;
    const int vi = e % vn;
    const int t0 = idxs[e * 2], t1 = idxs[e * 2 + 1];
    float x = 0.f, y = 0.f, z = 0.f;
    if ((unsigned)t0 < (unsigned)tn && (unsigned)t1 < (unsigned)tn) {
        const float2 d0 = *(const float2 *)(direct + ((long long)t0 * vn + vi) * 2);
        const float2 d1 = *(const float2 *)(direct + ((long long)t1 * vn + vi) * 2);
        const float2 c0 = *(const float2 *)(coords + (long long)t0 * 2);
        const float2 c1 = *(const float2 *)(coords + (long long)t1 * 2);
        const float lz1 = __fmaf_rn(d1.x, c1.y, -__fmul_rn(d1.y, c1.x));
        const float lz0 = __fmaf_rn(d0.x, c0.y, -__fmul_rn(d0.y, c0.x));
        z = __fmaf_rn(d0.x, d1.y, -__fmul_rn(d0.y, d1.x));
        x = __fmaf_rn(d1.x, lz0, -__fmul_rn(d0.x, lz1));
        y = __fmaf_rn(d1.y, lz0, -__fmul_rn(d0.y, lz1));
        const float vx0 = __fmul_rn(d0.x, __fmaf_rn(-c0.x, z, x)), vx1 = __fmul_rn(d1.x, __fmaf_rn(-c1.x, z, x));
        const float vy0 = __fmul_rn(d0.y, __fmaf_rn(-c0.y, z, y)), vy1 = __fmul_rn(d1.y, __fmaf_rn(-c1.y, z, y));
        if (vx0 < 0.f && vx1 < 0.f && vy0 < 0.f && vy1 < 0.f) { x = -x; y = -y; z = -z; }  // flip direction (:219-220)
        if (__fmul_rn(vx0, vx1) < 0.f || __fmul_rn(vy0, vy1) < 0.f) { x = 0.f; y = 0.f; z = 0.f; }  // rays do not meet (:222-223)
    }
    hyp[e * 3] = x;
    hyp[e * 3 + 1] = y;
    hyp[e * 3 + 2] = z;
}

// the reference's test (:285-300): diff = h.xy - c*h.z; both norms >= 1e-6 (double compare), both products
// direct*diff >= 0, |cos| > thresh with cos = rn(rn(dx*diffx) + rn(dy*diffy)) / rn(norm2*norm1)
__device__ __forceinline__ bool vp_test(float dx, float dy, float cx, float cy, float hx, float hy, float hz, float thresh)
{
    const float n1 = __fsqrt_rn(__fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
    const float fx = __fmaf_rn(-cx, hz, hx), fy = __fmaf_rn(-cy, hz, hy);
    const float n2 = __fsqrt_rn(__fmaf_rn(fx, fx, __fmul_rn(fy, fy)));
    if ((double)n1 < 1e-6 || (double)n2 < 1e-6) return false;
    const float vx = __fmul_rn(dx, fx), vy = __fmul_rn(dy, fy);
    if (vx < 0.f || vy < 0.f) return false;
    return fabsf(__fdiv_rn(__fadd_rn(vx, vy), __fmul_rn(n2, n1))) > thresh;
}

__global__ void __launch_bounds__(256) k_prim_vote_vp(const float *__restrict__ direct, const float *__restrict__ coords,
                                                      const float *__restrict__ hyp, uint8_t *__restrict__ inliers, int tn,
                                                      int vn, int hn, float thresh)
{
    const int rowi = blockIdx.y;  // hi*vn + vi
    const int vi = rowi % vn;
    const float hx = hyp[rowi * 3], hy = hyp[rowi * 3 + 1], hz = hyp[rowi * 3 + 2];
    for (int ti = blockIdx.x * blockDim.x + threadIdx.x; ti < tn; ti += gridDim.x * blockDim.x) {
        const float2 dir = *(const float2 *)(direct + ((long long)ti * vn + vi) * 2);
        const float2 c = *(const float2 *)(coords + (long long)ti * 2);
        if (vp_test(dir.x, dir.y, c.x, c.y, hx, hy, hz, thresh)) inliers[(long long)rowi * tn + ti] = 1;
    }
}

}  // namespace

extern "C" int pvv_generate_hypothesis_vanishing_point(const float *direct, const float *coords, const int32_t *idxs, int tn,
                                                       int vn, int hn, float *hyp, void *stream)
{
    if (!direct || !coords || !idxs || !hyp || tn < 0 || vn <= 0 || hn < 0) return PVV_ERR_INVALID;
    if (hn == 0) return PVV_OK;
    k_prim_generate_vp<<<(hn * vn + 255) / 256, 256, 0, (cudaStream_t)stream>>>(direct, coords, idxs, tn, vn, hn, hyp);
    return cudaGetLastError() == cudaSuccess ? PVV_OK : PVV_ERR_CUDA;
}

extern "C" int pvv_voting_for_hypothesis_vanishing_point(const float *direct, const float *coords, const float *hyp,
                                                         uint8_t *inliers, int tn, int vn, int hn, float inlier_thresh,
                                                         void *stream)
{
    if (!direct || !coords || !hyp || !inliers || tn < 0 || vn <= 0 || hn < 0) return PVV_ERR_INVALID;
    if (hn == 0 || tn == 0) return PVV_OK;
    if (vn > 65535) return PVV_ERR_UNSUPPORTED;
    const int step = (65535 / vn) * vn;
    for (long long r0 = 0; r0 < (long long)hn * vn; r0 += step) {
        const int rows = (int)min((long long)step, (long long)hn * vn - r0);
        const dim3 grid(min((tn + 255) / 256, 64), rows);
        k_prim_vote_vp<<<grid, 256, 0, (cudaStream_t)stream>>>(direct, coords, hyp + r0 * 3, inliers + r0 * tn, tn, vn, hn,
                                                               inlier_thresh);
    }
    return cudaGetLastError() == cudaSuccess ? PVV_OK : PVV_ERR_CUDA;
}

extern "C" int pvv_generate_hypothesis(const float *direct, const float *coords, const int32_t *idxs, int tn, int vn,
                                       int hn, float *hyp, void *stream)
{
    if (!direct || !coords || !idxs || !hyp || tn < 0 || vn <= 0 || hn < 0) return PVV_ERR_INVALID;
    if (hn == 0) return PVV_OK;
    k_prim_generate<<<(hn * vn + 255) / 256, 256, 0, (cudaStream_t)stream>>>(direct, coords, idxs, tn, vn, hn, hyp);
    return cudaGetLastError() == cudaSuccess ? PVV_OK : PVV_ERR_CUDA;
}

extern "C" int pvv_voting_for_hypothesis(const float *direct, const float *coords, const float *hyp, uint8_t *inliers,
                                         int tn, int vn, int hn, float inlier_thresh, void *stream)
{
    if (!direct || !coords || !hyp || !inliers || tn < 0 || vn <= 0 || hn < 0) return PVV_ERR_INVALID;
    if (hn == 0 || tn == 0) return PVV_OK;
    if (vn > 65535) return PVV_ERR_UNSUPPORTED;
    // grid.y is limited to 65535 rows per launch: step by whole hypotheses so that row % vn stays the keypoint
    const int step = (65535 / vn) * vn;
    for (long long r0 = 0; r0 < (long long)hn * vn; r0 += step) {
        const int rows = (int)min((long long)step, (long long)hn * vn - r0);
        const dim3 grid(min((tn + 255) / 256, 64), rows);
        k_prim_vote<<<grid, 256, 0, (cudaStream_t)stream>>>(direct, coords, hyp + r0 * 2, inliers + r0 * tn, tn, vn, hn,
                                                            inlier_thresh);
    }
    return cudaGetLastError() == cudaSuccess ? PVV_OK : PVV_ERR_CUDA;
}

extern "C" int pvv_vote_counts(const float *direct, const float *coords, const float *hyp, int tn, int vn, int hn,
                               float inlier_thresh, int32_t *counts, void *stream)
{
    if (!direct || !coords || !hyp || !counts || tn < 0 || vn <= 0 || hn < 0) return PVV_ERR_INVALID;
    if (hn == 0) return PVV_OK;
    k_prim_counts<<<hn * vn, 256, 0, (cudaStream_t)stream>>>(direct, coords, hyp, tn, vn, hn, inlier_thresh, counts);
    return cudaGetLastError() == cudaSuccess ? PVV_OK : PVV_ERR_CUDA;
}
