// pvv_common.cuh -- shared device helpers for libpvvote (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/pvvote.h"

#define PVV_MAX_VN 64
#define PVV_TILE_PX 1024   // pixels per compaction tile (K1)
#define PVV_CHUNK_PX 256   // pixels per voting work item (K3)
#define PVV_POOL_ALIGN 16  // every image starts on a multiple of this many pixels in the pool: the SoA rows of a voting
                           // chunk (4 / 8 / 16 bytes per pixel) are then 16-byte aligned sources for bulk copies (TMA)
// Checked build (-DPVV_CHECKED, tools/build_variant.sh checked -DPVV_CHECKED): bounds asserts on pool / chunk-table / queue /
// tensor-memory indices.  A failed check prints the site and traps (the call returns PVV_ERR_CUDA).  Off in the shipped
// library: the macro expands to nothing.  This replaces a compute-sanitizer run, which the GPU pool does not allow.
#ifdef PVV_CHECKED
#include <stdio.h>
#define PVV_CHECK(cond)                                                                                              \
    do {                                                                                                             \
        if (!(cond)) {                                                                                               \
            printf("PVV_CHECK failed: %s  (%s:%d, block %d thread %d)\n", #cond, __FILE__, __LINE__, (int)blockIdx.x, \
                   (int)threadIdx.x);                                                                                \
            __trap();                                                                                                \
        }                                                                                                            \
    } while (0)
#else
#define PVV_CHECK(cond) ((void)0)
#endif

#ifndef PVV_FIN_SPLIT
#define PVV_FIN_SPLIT 4     // CTAs sharing the pixels of one (image, keypoint) in K4
#endif

// ---------------------------------------------------------------------------------------------------
// Reference arithmetic (bit-exact restatement of what nvcc 12.9 emits for the reference kernels on
// sm_100a; see DESIGN.md "Reference arithmetic").  Explicit _rn intrinsics: nothing may be re-contracted.
// ---------------------------------------------------------------------------------------------------

// ||n|| of a direction vector: sqrt_rn(fma(nx,nx, rn(ny*ny)))   (ransac_voting_kernel.cu:118)
__device__ __forceinline__ float pvv_norm1(float nx, float ny)
{
    return __fsqrt_rn(__fmaf_rn(nx, nx, __fmul_rn(ny, ny)));
}

// One (hypothesis, keypoint, pixel) test with the reference's rounding sequence
// (ransac_voting_kernel.cu:112-125).  n1 = pvv_norm1(nx, ny) may be hoisted by the caller.
__device__ __forceinline__ bool pvv_exact_test(float nx, float ny, float n1, float cx, float cy, float hx, float hy,
                                               float thresh)
{
    const float dx = __fsub_rn(hx, cx);
    const float dy = __fsub_rn(hy, cy);
    const float n2 = __fsqrt_rn(__fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
    if ((double)n1 < 1e-6 || (double)n2 < 1e-6) return false;
    const float den = __fmul_rn(n2, n1);
    const float num = __fmaf_rn(nx, dx, __fmul_rn(ny, dy));
    return __fdiv_rn(num, den) > thresh;
}

// 2-line intersection (ransac_voting_kernel.cu:27-48).  Returns false (hypothesis stays 0,0) for
// degenerate pairs.  d0/d1 are the raw direction vectors of the two pixels, c0/c1 their coordinates.
__device__ __forceinline__ bool pvv_intersect(float d0x, float d0y, float cx0, float cy0, float d1x, float d1y,
                                              float cx1, float cy1, float &x, float &y)
{
    const float p = __fmul_rn(d0y, d1x);
    const float q = __fmul_rn(d0x, d1y);
    const float det_y = __fsub_rn(p, q);
    if (fabs((double)det_y) < 1e-6) return false;
    const float det_x = __fsub_rn(q, p);
    if (fabs((double)det_x) < 1e-6) return false;
    const float s1 = __fmaf_rn(d1y, cx1, -__fmul_rn(d1x, cy1));
    const float s0 = __fmaf_rn(d0y, cx0, -__fmul_rn(d0x, cy0));
    const float num_y = __fmaf_rn(d1y, s0, -__fmul_rn(d0y, s1));
    const float num_x = __fmaf_rn(d0x, s1, -__fmul_rn(d1x, s0));
    x = __fdiv_rn(num_x, det_x);
    y = __fdiv_rn(num_y, det_y);
    return true;
}

// ---------------------------------------------------------------------------------------------------
// Philox4x32-10 counter-based generator: the library's own, seedable, shard-invariant draw stream.
// ---------------------------------------------------------------------------------------------------
struct pvv_u4 { uint32_t x, y, z, w; };

__host__ __device__ __forceinline__ pvv_u4 pvv_philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                      uint32_t k0, uint32_t k1)
{
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint64_t m0 = (uint64_t)0xD2511F53u * c0;
        const uint64_t m1 = (uint64_t)0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t)(m1 >> 32) ^ c1 ^ k0;
        const uint32_t n1 = (uint32_t)m1;
        const uint32_t n2 = (uint32_t)(m0 >> 32) ^ c3 ^ k1;
        const uint32_t n3 = (uint32_t)m0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    pvv_u4 o; o.x = c0; o.y = c1; o.z = c2; o.w = c3;
    return o;
}

#define PVV_STREAM_IDXS 0x1D5u
#define PVV_STREAM_SELECT 0x5E1u

// pixel-pair indices for (image, hypothesis hi, keypoint vi): two u32 draws reduced mod tn
__device__ __forceinline__ void pvv_draw_pair(uint64_t seed, int64_t image_id, int hi, int vi, int vn, int tn,
                                              int &t0, int &t1)
{
    const uint64_t e = (uint64_t)hi * (uint64_t)vn + (uint64_t)vi;
    const pvv_u4 r = pvv_philox((uint32_t)e, (uint32_t)(e >> 32) ^ (PVV_STREAM_IDXS << 16), (uint32_t)image_id,
                                (uint32_t)((uint64_t)image_id >> 32), (uint32_t)seed, (uint32_t)(seed >> 32));
    t0 = (int)(r.x % (uint32_t)tn);
    t1 = (int)(r.y % (uint32_t)tn);
}

// uniform in [0,1) for (image, pixel): top 24 bits
__device__ __forceinline__ float pvv_draw_uniform(uint64_t seed, int64_t image_id, uint32_t pixel)
{
    const pvv_u4 r = pvv_philox(pixel, (PVV_STREAM_SELECT << 16), (uint32_t)image_id,
                                (uint32_t)((uint64_t)image_id >> 32), (uint32_t)seed, (uint32_t)(seed >> 32));
    return (float)(r.x >> 8) * 5.9604644775390625e-8f;  // 2^-24
}

// `max_num / fg.float()` in the reference is Tensor.__rtruediv__: reciprocal(fg) * max_num, in fp32
__device__ __forceinline__ float pvv_subsample_prob(int max_num, long long fg)
{
    return __fmul_rn(__frcp_rn((float)fg), (float)max_num);
}

// ---------------------------------------------------------------------------------------------------
// mask access
// ---------------------------------------------------------------------------------------------------
// Returns the reference's `.byte()` value of the element (PVV_FG_BYTE) or 1/0 for `== 1` (PVV_FG_EQ1).
__device__ __forceinline__ uint32_t pvv_mask_value(const void *mask, int dtype, int64_t off, int fg_rule)
{
    long long v;
    switch (dtype) {
        case PVV_MASK_U8: v = ((const uint8_t *)mask)[off]; break;
        case PVV_MASK_I32: v = ((const int32_t *)mask)[off]; break;
        case PVV_MASK_I64: v = ((const long long *)mask)[off]; break;
        default: {  // float: .byte() truncates toward zero, == 1 compares the float
            const float f = ((const float *)mask)[off];
            if (fg_rule == PVV_FG_EQ1) return f == 1.0f ? 1u : 0u;
            v = (long long)f;
        }
    }
    if (fg_rule == PVV_FG_EQ1) return v == 1 ? 1u : 0u;
    return (uint32_t)(v & 0xFF);
}

// pool pixels an image of n compacted pixels occupies (its start is aligned, see PVV_POOL_ALIGN)
__host__ __device__ __forceinline__ int pvv_pool_span(int n) { return (n + PVV_POOL_ALIGN - 1) & ~(PVV_POOL_ALIGN - 1); }

struct pvv_ws;  // workspace view, pvv_capi.cu
