// pvv_vote.cu -- K2 (hypothesis generation) and K3 (fused inlier counting, the FP32-bound hot loop).
//
// K2 replaces generate_hypothesis (ransac_voting_kernel.cu:11-86) for the whole batch.
// K3 replaces  zeros([hn,vn,tn]) + voting_for_hypothesis (ransac_voting_kernel.cu:88-167) + torch.sum(inlier, 2)
//              (ransac_voting_gpu.py:563-567 / :371-374): the [hn,vn,tn] uint8 matrix is never materialised.
//
// K3 design (DESIGN.md "K3"):
//   * persistent CTAs, dynamic work queue of (pixel chunk, keypoint) items;
//   * the chunk's pixels are staged in shared memory as 6 derived floats per pixel, read back as warp-broadcast
//     LDS.128 + LDS.64; every thread keeps H hypotheses and H integer counters in registers, so the inner loop has
//     no atomics, no shuffles and no global traffic;
//   * decision by a guard-banded filter in a rotated frame:  inlier  <=>  a > tau*|c|  with a = d.n^, c = d x n^,
//     tau = t/sqrt(1-t^2): 4 FFMA + 2 FSETP per test.  Tests that fall inside the band (|a-|c'|| <= eps_rel*a+eps_abs)
//     are re-evaluated with the reference's exact rounding sequence (pvv_exact_test), which makes the counts
//     bit-identical to the reference kernel's (error analysis in DESIGN.md "Guard band").
#include <stdlib.h>

#include "pvv_internal.cuh"

namespace {

// ---- K2 -------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_hypotheses(pvv_dims d, const int32_t *idxs, pvv_ws ws, float *hyp_out)
{
    const int img = blockIdx.y;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= d.vn * d.hn) return;
    const int vi = e / d.hn, hi = e - vi * d.hn;
    const int tn = ws.tn[img];
    float2 out = make_float2(0.f, 0.f);
    int t0 = 0, t1 = 0;
    if (tn > 0) {
        if (idxs) {
            const long long o = (((long long)img * d.hn + hi) * d.vn + vi) * 2;
            t0 = idxs[o];
            t1 = idxs[o + 1];
        } else {
            pvv_draw_pair(d.seed, d.image_id_base + img, hi, vi, d.vn, tn, t0, t1);
        }
        if ((unsigned)t0 < (unsigned)tn && (unsigned)t1 < (unsigned)tn) {
            const long long off = ws.img_off[img];
            const uint32_t xy0 = ws.pix_xy[off + t0], xy1 = ws.pix_xy[off + t1];
            const float2 d0 = ws.pix_dir[(long long)vi * ws.cap + off + t0];
            const float2 d1 = ws.pix_dir[(long long)vi * ws.cap + off + t1];
            float x, y;
            if (pvv_intersect(d0.x, d0.y, (float)(xy0 & 0xffffu), (float)(xy0 >> 16), d1.x, d1.y,
                              (float)(xy1 & 0xffffu), (float)(xy1 >> 16), x, y))
                out = make_float2(x, y);
        }
    }
    ws.hyp[((long long)img * d.vn + vi) * ws.hn_pad + hi] = out;
    const long long ro = (((long long)img * d.hn + hi) * d.vn + vi) * 2;
    if (hyp_out) { hyp_out[ro] = out.x; hyp_out[ro + 1] = out.y; }
    if (ws.idxs_used) { ws.idxs_used[ro] = t0; ws.idxs_used[ro + 1] = t1; }
}

// ---- K3 -------------------------------------------------------------------------------------------
struct band_consts {
    float tau;        // t / sqrt(1 - t^2)
    float krel;       // half-width of the band per unit of max |coordinate| (covers tau's uncertainty + rounding)
    float abs_floor;  // catches ||d|| < 1e-6 (reference skips those)
};

constexpr int K3_GROUP = 4;    // pixels per band-check group in the hot loop (two packed pixel pairs)
constexpr int K3_QCAP = 1024;  // deferred (thread, pixel group) entries per work item; overflow is resolved in line
constexpr int K3_PAIRS = PVV_CHUNK_PX / 2;

// packed FP32x2 helpers: Blackwell issues one FFMA2 for two IEEE fp32 FMAs (half the issue slots of two FFMAs)
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk(float lo, float hi)
{
    f32x2 d;
    asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(lo), "f"(hi));
    return d;
}
__device__ __forceinline__ void unpk(f32x2 v, float &lo, float &hi)
{
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 fadd2(f32x2 a, f32x2 b)
{
    f32x2 d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ f32x2 ffma2(f32x2 a, f32x2 b, f32x2 c)
{
    f32x2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}

// Per-pixel constants of the filter, in the chunk-local frame (origin = first pixel of the chunk):
//   a = hx*ax + hy*ay + az   (= d . n^)          c' = hx*cx + hy*cy + cz   (= tau * d x n^)
// Shared-memory layout: pixel PAIRS, 12 floats (3 x LDS.128) per pair, every constant as {pixel 2q, pixel 2q+1} so
// that one FFMA2 evaluates a term for two pixels:  [ax ax' ay ay'] [az az' cx cx'] [cy cy' cz cz'].
struct px_consts { float ax, ay, az, cx, cy, cz; };

__device__ __forceinline__ px_consts k3_load_px(const float *sP, int j)
{
    const float *q = sP + (j >> 1) * 12 + (j & 1);
    px_consts c;
    c.ax = q[0]; c.ay = q[2]; c.az = q[4]; c.cx = q[6]; c.cy = q[8]; c.cz = q[10];
    return c;
}
__device__ __forceinline__ void k3_store_px(float *sP, int j, const px_consts &c)
{
    float *q = sP + (j >> 1) * 12 + (j & 1);
    q[0] = c.ax; q[2] = c.ay; q[4] = c.az; q[6] = c.cx; q[8] = c.cy; q[10] = c.cz;
}

// Hypothesis in the chunk-local frame and its band half-width T (constant over the chunk: |a| <= |hx|+|hy|+P).
// Used verbatim by the hot loop's set-up and by the deferred pass, so both see bit-identical band tests.
__device__ __forceinline__ void k3_local_hyp(float hxr, float hyr, int ox, int oy, float P, const band_consts &bc,
                                             float &hx, float &hy, float &T)
{
    hx = __fsub_rn(hxr, (float)ox);
    hy = __fsub_rn(hyr, (float)oy);
    const float m = __fadd_rn(__fadd_rn(fabsf(hx), fabsf(hy)), P);
    if (m < 1e15f) {
        T = __fmaf_rn(bc.krel, m, bc.abs_floor);
    } else {
        // far-away / non-finite hypothesis: every test takes the exact path (the reference's overflow behaviour);
        // finite stand-ins keep NaN out of the sign-bit counting
        hx = 0.f; hy = 0.f; T = __int_as_float(0x7f800000);
    }
}

// e = a - |c'| for one test, scalar form (the hot loop computes the same two FMA chains with FFMA2: identical bits)
__device__ __forceinline__ float k3_e(float hx, float hy, const px_consts &c)
{
    const float a = __fmaf_rn(hx, c.ax, __fmaf_rn(hy, c.ay, c.az));
    const float cc = __fmaf_rn(hx, c.cx, __fmaf_rn(hy, c.cy, c.cz));
    return __fsub_rn(a, fabsf(cc));
}

// Re-evaluation of ONE deferred test: hypothesis slot k of thread `t` at pixel `j`.  The hot loop counted every test
// inside the guard band as "not inlier"; here those get the reference's exact decision, added straight to the global
// counts.  The band test is recomputed from the same staged constants (sP, sH) with the same roundings as the hot loop.
template <int THREADS>
__device__ __forceinline__ void k3_resolve_one(const pvv_dims &d, const pvv_ws &ws, long long row, int hbase, int t, int j,
                                               int k, int n, const float *sP, const float4 *sH, const float2 *sDir,
                                               const uint32_t *sXY, unsigned long long *amb)
{
    const int h = hbase + k * THREADS + t;
    if (j >= n || h >= d.hn) return;
    const float4 hl = sH[k * THREADS + t];  // {hx, hy, T} in the chunk-local frame
    const float e = k3_e(hl.x, hl.y, k3_load_px(sP, j));
    if (fabsf(e) > hl.z) return;
    const float2 hv = ws.hyp[row + h];
    const float2 dir = sDir[j];
    const uint32_t xy = sXY[j];
    if (pvv_exact_test(dir.x, dir.y, pvv_norm1(dir.x, dir.y), (float)(xy & 0xffffu), (float)(xy >> 16), hv.x, hv.y, d.thresh))
        atomicAdd(&ws.counts[row + h], 1);
    ++*amb;
}

// work item -> (image, keypoint, pixel range)
struct k3_item { int img, vi, p0, n, ox, oy; long long off; };

__device__ __forceinline__ k3_item k3_decode_item(const pvv_dims &d, const pvv_ws &ws, int item)
{
    k3_item it;
    const int ci = item / d.vn;
    it.vi = item - ci * d.vn;
    int lo = 0, hi = d.b;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (ws.chunk_start[mid] <= ci) lo = mid; else hi = mid;
    }
    it.img = lo;
    const int tn = ws.tn[lo];
    it.off = ws.img_off[lo];
    it.p0 = (ci - ws.chunk_start[lo]) * PVV_CHUNK_PX;
    it.n = min(PVV_CHUNK_PX, tn - it.p0);
    const uint32_t xy0 = ws.pix_xy[it.off + it.p0];
    it.ox = (int)(xy0 & 0xffffu);
    it.oy = (int)(xy0 >> 16);
    return it;
}

// exact-only kernel: every test with the reference's rounding sequence (force_exact, thresholds outside the filter range)
template <int H, int THREADS>
__global__ void __launch_bounds__(THREADS) k_count_exact(pvv_dims d, pvv_ws ws)
{
    __shared__ float2 sDir[PVV_CHUNK_PX];
    __shared__ uint32_t sXY[PVV_CHUNK_PX];
    __shared__ int s_item;
    const int tid = threadIdx.x;
    const int total_items = ws.chunk_start[d.b] * d.vn;
    while (true) {
        __syncthreads();
        if (tid == 0) s_item = (int)atomicAdd(&ws.counters[PVV_CTR_WORK], 1u);
        __syncthreads();
        const int item = s_item;
        if (item >= total_items) break;
        const k3_item it = k3_decode_item(d, ws, item);
        for (int j = tid; j < it.n; j += THREADS) {
            sXY[j] = ws.pix_xy[it.off + it.p0 + j];
            sDir[j] = ws.pix_dir[(long long)it.vi * ws.cap + it.off + it.p0 + j];
        }
        __syncthreads();
        const long long row = ((long long)it.img * d.vn + it.vi) * ws.hn_pad;
        for (int hbase = 0; hbase + (tid & ~31) < d.hn; hbase += THREADS * H) {
            float hxr[H], hyr[H];
            int cnt[H];
#pragma unroll
            for (int k = 0; k < H; ++k) {
                const int h = hbase + k * THREADS + tid;
                const float2 hv = h < d.hn ? ws.hyp[row + h] : make_float2(0.f, 0.f);
                hxr[k] = hv.x; hyr[k] = hv.y; cnt[k] = 0;
            }
            for (int j = 0; j < it.n; ++j) {
                const uint32_t xy = sXY[j];
                const float2 q = sDir[j];
                const float cx = (float)(xy & 0xffffu), cy = (float)(xy >> 16);
                const float n1 = pvv_norm1(q.x, q.y);
#pragma unroll
                for (int k = 0; k < H; ++k) cnt[k] += pvv_exact_test(q.x, q.y, n1, cx, cy, hxr[k], hyr[k], d.thresh);
            }
#pragma unroll
            for (int k = 0; k < H; ++k) {
                const int h = hbase + k * THREADS + tid;
                if (h < d.hn && cnt[k]) atomicAdd(&ws.counts[row + h], cnt[k]);
            }
        }
    }
}

// guard-banded filter kernel (MODE: code-generation variants of the same arithmetic, see the constants below)
template <int H, int THREADS, int MODE>
__global__ void __launch_bounds__(THREADS) k_count(pvv_dims d, pvv_ws ws, band_consts bc)
{
    constexpr bool MIN3 = (MODE & 1) != 0;   // band check by FMNMX3 min-tracking instead of an FSETP chain
    constexpr bool PACKG = (MODE & 2) != 0;  // T - e for a pixel pair as one FADD2
    __shared__ __align__(16) float sP[12 * K3_PAIRS];
    __shared__ float2 sDir[PVV_CHUNK_PX];
    __shared__ uint32_t sXY[PVV_CHUNK_PX];
    __shared__ float4 sH[H * THREADS];  // this pass's hypotheses in the chunk-local frame: {hx, hy, T, -}
    __shared__ uint32_t s_q[K3_QCAP];
    __shared__ int s_item, s_P, s_qn;

    const int tid = threadIdx.x;
    const int total_items = ws.chunk_start[d.b] * d.vn;
    unsigned long long amb_local = 0;
    const px_consts never = {0.f, 0.f, -1e30f, 0.f, 0.f, 0.f};  // a = -1e30: never an inlier, never in the band

    while (true) {
        __syncthreads();  // previous item's smem fully consumed
        if (tid == 0) { s_item = (int)atomicAdd(&ws.counters[PVV_CTR_WORK], 1u); s_P = 0; }
        __syncthreads();
        const int item = s_item;
        if (item >= total_items) break;
        const k3_item it = k3_decode_item(d, ws, item);
        const int n = it.n, ox = it.ox, oy = it.oy;

        // ---- stage the chunk: raw values + derived constants in a chunk-local frame ----
        const int npad = (n + K3_GROUP - 1) / K3_GROUP * K3_GROUP;
        for (int j = tid; j < npad; j += THREADS) {
            px_consts c = never;
            if (j < n) {
                const uint32_t xy = ws.pix_xy[it.off + it.p0 + j];
                const float2 dir = ws.pix_dir[(long long)it.vi * ws.cap + it.off + it.p0 + j];
                sXY[j] = xy;
                sDir[j] = dir;
                const float n1 = pvv_norm1(dir.x, dir.y);
                // the reference skips ||n|| < 1e-6; a non-finite norm degenerates its cosine: never an inlier either
                if ((double)n1 >= 1e-6 && n1 < 3.0e38f) {
                    const double inv = rsqrt((double)dir.x * dir.x + (double)dir.y * dir.y);
                    const int px = (int)(xy & 0xffffu) - ox, py = (int)(xy >> 16) - oy;
                    c.ax = (float)(dir.x * inv);
                    c.ay = (float)(dir.y * inv);
                    c.cx = (float)((double)bc.tau * (dir.y * inv));
                    c.cy = (float)(-(double)bc.tau * (dir.x * inv));
                    c.az = (float)(-((double)px * c.ax + (double)py * c.ay));
                    c.cz = (float)(-((double)px * c.cx + (double)py * c.cy));
                    atomicMax(&s_P, abs(px) + abs(py));
                }
            }
            k3_store_px(sP, j, c);
        }
        __syncthreads();
        const float P = (float)s_P;
        const long long row = ((long long)it.img * d.vn + it.vi) * ws.hn_pad;

        for (int hbase = 0; hbase < d.hn; hbase += THREADS * H) {
            const bool active = hbase + (tid & ~31) < d.hn;  // warp-uniform: does this warp own hypotheses in the pass
            if (tid == 0) s_qn = 0;
            __syncthreads();  // also: the previous pass's deferred entries (s_q, sH) are consumed
            if (active) {
                f32x2 hx2[H], hy2[H];
                float T[H];
                int cnt[H];
#pragma unroll
                for (int k = 0; k < H; ++k) {
                    const int h = hbase + k * THREADS + tid;
                    const float2 hv = h < d.hn ? ws.hyp[row + h] : make_float2(0.f, 0.f);
                    float hx, hy;
                    k3_local_hyp(hv.x, hv.y, ox, oy, P, bc, hx, hy, T[k]);
                    hx2[k] = pk(hx, hx);
                    hy2[k] = pk(hy, hy);
                    cnt[k] = 0;
                    sH[k * THREADS + tid] = make_float4(hx, hy, T[k], 0.f);  // for the deferred pass
                }
                // Groups of K3_GROUP pixels (two packed pairs) share one band check.  A group with any test inside the
                // guard band only sets a bit in a per-thread mask: the hot loop has no branch and no divergence.
                uint32_t dmask[2] = {0u, 0u};
                const ulonglong2 *pp = reinterpret_cast<const ulonglong2 *>(sP);
                const int ngroups = npad / K3_GROUP;
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    const int g_end = min(ngroups, 32 * (half + 1));
                    uint32_t bit = 1u, dm = 0u;
                    for (int g = 32 * half; g < g_end; ++g, pp += 3 * (K3_GROUP / 2), bit <<= 1) {
                        bool clear = true;
                        float mabs[H];
                        if (MIN3) {
#pragma unroll
                            for (int k = 0; k < H; ++k) mabs[k] = __int_as_float(0x7f800000);
                        }
#pragma unroll
                        for (int u = 0; u < K3_GROUP / 2; ++u) {
                            const ulonglong2 v0 = pp[3 * u], v1 = pp[3 * u + 1], v2 = pp[3 * u + 2];
#pragma unroll
                            for (int k = 0; k < H; ++k) {
                                const f32x2 a2 = ffma2(hx2[k], v0.x, ffma2(hy2[k], v0.y, v1.x));
                                const f32x2 c2 = ffma2(hx2[k], v1.y, ffma2(hy2[k], v2.x, v2.y));
                                float a0, a1, c0, c1;
                                unpk(a2, a0, a1);
                                unpk(c2, c0, c1);
                                // ne = |c'| - a = -e.  T + ne < 0  <=>  e > T: definitely an inlier.  NaN cannot reach
                                // here (k3_local_hyp).
                                const float ne0 = __fsub_rn(fabsf(c0), a0), ne1 = __fsub_rn(fabsf(c1), a1);
                                if (PACKG) {
                                    float g0, g1;
                                    unpk(fadd2(pk(T[k], T[k]), pk(ne0, ne1)), g0, g1);
                                    cnt[k] += (int)(__float_as_uint(g0) >> 31);
                                    cnt[k] += (int)(__float_as_uint(g1) >> 31);
                                } else {
                                    cnt[k] += (int)(__float_as_uint(__fadd_rn(T[k], ne0)) >> 31);
                                    cnt[k] += (int)(__float_as_uint(__fadd_rn(T[k], ne1)) >> 31);
                                }
                                if (MIN3) mabs[k] = fminf(fminf(mabs[k], fabsf(ne0)), fabsf(ne1));
                                else clear = clear && (fabsf(ne0) > T[k]) && (fabsf(ne1) > T[k]);
                            }
                        }
                        if (MIN3) {
#pragma unroll
                            for (int k = 0; k < H; ++k) clear = clear && (mabs[k] > T[k]);
                        }
                        if (!clear) dm |= bit;
                    }
                    dmask[half] = dm;
                }
#pragma unroll
                for (int k = 0; k < H; ++k) {
                    const int h = hbase + k * THREADS + tid;
                    if (h < d.hn && cnt[k]) atomicAdd(&ws.counts[row + h], cnt[k]);
                }
                // queue this thread's deferred groups for the dense exact pass
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    uint32_t dm = dmask[half];
                    while (dm) {
                        const int grp = 32 * half + __ffs(dm) - 1;
                        dm &= dm - 1;
                        const int q = atomicAdd(&s_qn, 1);
                        if (q < K3_QCAP) {
                            s_q[q] = ((uint32_t)grp << 16) | (uint32_t)tid;
                        } else {  // queue full (pathological inputs): resolve in line
                            for (int sub = 0; sub < K3_GROUP * H; ++sub)
                                k3_resolve_one<THREADS>(d, ws, row, hbase, tid, grp * K3_GROUP + (sub % K3_GROUP), sub / K3_GROUP,
                                                        n, sP, sH, sDir, sXY, &amb_local);
                        }
                    }
                }
            }
            __syncthreads();
            // deferred pass: one (pixel, hypothesis) re-evaluation per thread and step, spread over the whole CTA
            const int qtot = min(s_qn, K3_QCAP) * (K3_GROUP * H);
            for (int idx = tid; idx < qtot; idx += THREADS) {
                const uint32_t ent = s_q[idx / (K3_GROUP * H)];
                const int sub = idx % (K3_GROUP * H);
                k3_resolve_one<THREADS>(d, ws, row, hbase, (int)(ent & 0xffffu), (int)(ent >> 16) * K3_GROUP + (sub % K3_GROUP),
                                        sub / K3_GROUP, n, sP, sH, sDir, sXY, &amb_local);
            }
        }
    }
    if (amb_local) atomicAdd((unsigned long long *)&ws.counters[PVV_CTR_AMB_LO], amb_local);
}
}  // namespace

cudaError_t pvv_launch_hypotheses(const pvv_dims &d, const int32_t *idxs, const pvv_ws &ws, float *hyp_out,
                                  cudaStream_t st, int *launches)
{
    const dim3 grid((d.vn * d.hn + 255) / 256, d.b);
    k_hypotheses<<<grid, 256, 0, st>>>(d, idxs, ws, hyp_out);
    ++*launches;
    return cudaGetLastError();
}

// Guard-band constants for threshold t (DESIGN.md "Guard band").  Returns false when the fast filter does
// not apply (t outside [0.5, 0.99999]): the caller then runs the exact-only kernel.
static bool pvv_band_for(float t, band_consts *bc)
{
    if (!(t >= 0.5f && t <= 0.99999f)) return false;
    const double u = 5.9604644775390625e-8;  // 2^-24
    const double td = (double)t;
    const double one_m = 1.0 - td * td;
    const double tau = td / sqrt(one_m);
    // first-order bound on the relative rounding error of the reference's quotient num/(n1*n2):
    // n1, n2: 2u each (fma + sqrt), den: +u, num: (1 + 1/t)u (no cancellation near the threshold), div: +u
    const double eq = (7.0 + 1.0 / td) * u;
    const double etau = eq / one_m;                 // induced relative uncertainty of tau
    if (etau > 0.05) return false;
    const double erel = 1.10 * etau / (1.0 - etau) + 8.0 * u;      // uncertainty of the reference's own decision
    const double kabs = 12.0 * u * (1.0 + tau) * (1.0 + etau);     // rounding of the filter's a and c' (<= 6.5u(1+tau)m)
    bc->tau = (float)tau;
    bc->krel = (float)((erel + kabs) * (1.0 + 4.0 * u));
    bc->abs_floor = (float)(4e-6 * (1.0 + tau));
    return true;
}

template <int H, int THREADS, int MODE>
static cudaError_t launch_count_t(const pvv_dims &d, const pvv_ws &ws, const band_consts &bc, int sms, cudaStream_t st)
{
    int per_sm = 8;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_count<H, THREADS, MODE>, THREADS, 0);
    if (per_sm < 1) per_sm = 1;
    k_count<H, THREADS, MODE><<<sms * per_sm, THREADS, 0, st>>>(d, ws, bc);
    return cudaGetLastError();
}

cudaError_t pvv_launch_count(const pvv_dims &d, const pvv_ws &ws, int force_exact, cudaStream_t st, int *launches)
{
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    band_consts bc = {0.f, 0.f, 0.f};
    const bool fast = !force_exact && pvv_band_for(d.thresh, &bc);
    const char *venv = getenv("PVV_K3_VARIANT");  // tuning knob for experiments
    const int variant = venv ? atoi(venv) : 0;
    cudaError_t e;
    if (!fast) {
        int per_sm = 8;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_count_exact<4, 128>, 128, 0);
        k_count_exact<4, 128><<<sms * (per_sm < 1 ? 1 : per_sm), 128, 0, st>>>(d, ws);
        e = cudaGetLastError();
    }
    else if (variant == 1) e = launch_count_t<4, 128, 1>(d, ws, bc, sms, st);
    else if (variant == 2) e = launch_count_t<4, 128, 2>(d, ws, bc, sms, st);
    else if (variant == 3) e = launch_count_t<4, 128, 3>(d, ws, bc, sms, st);
    else if (variant == 4) e = launch_count_t<8, 64, 0>(d, ws, bc, sms, st);
    else if (variant == 5) e = launch_count_t<8, 64, 3>(d, ws, bc, sms, st);
    else if (variant == 6) e = launch_count_t<8, 128, 3>(d, ws, bc, sms, st);
    else if (variant == 7) e = launch_count_t<2, 256, 3>(d, ws, bc, sms, st);
    else e = launch_count_t<4, 128, 0>(d, ws, bc, sms, st);
    ++*launches;
    return e;
}
