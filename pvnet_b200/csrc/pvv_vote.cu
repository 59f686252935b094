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
    float erel;       // relative half-width of the band (in units of a)
    float kabs;       // absolute half-width per unit of coordinate magnitude
    float abs_floor;  // catches ||d|| < 1e-6 (reference skips those)
};

constexpr int K3_GROUP = 4;   // pixels per band-check group in the hot loop (PVV_CHUNK_PX is a multiple)
constexpr int K3_QCAP = 1024;  // deferred (thread, pixel) entries per work item; overflow is resolved in line

// Hypothesis in the chunk-local frame and its absolute band term.  Used verbatim by the hot loop's set-up and by
// the deferred pass, so both see bit-identical band tests (explicit _rn intrinsics: nothing may be re-contracted).
__device__ __forceinline__ void k3_local_hyp(float hxr, float hyr, int ox, int oy, float P, const band_consts &bc,
                                             float &hx, float &hy, float &eabs)
{
    hx = __fsub_rn(hxr, (float)ox);
    hy = __fsub_rn(hyr, (float)oy);
    const float m = __fadd_rn(__fadd_rn(fabsf(hx), fabsf(hy)), P);
    if (m < 1e15f) {
        eabs = __fmaf_rn(bc.kabs, m, bc.abs_floor);
    } else {
        // far-away / non-finite hypothesis: every test takes the exact path (the reference's overflow behaviour);
        // finite stand-ins keep NaN out of the sign-bit counting
        hx = 0.f; hy = 0.f; eabs = __int_as_float(0x7f800000);
    }
}

// e = a - |c'| and the band half-width t1 for one test; A = {n^x, n^y, -s, tau n^y}, Q = {-tau n^x, -tau s2, ., .}
__device__ __forceinline__ void k3_band(float hx, float hy, float eabs, float erel, const float4 &A, const float4 &Q,
                                        float &e, float &t1)
{
    const float a = __fmaf_rn(hx, A.x, __fmaf_rn(hy, A.y, A.z));
    const float c = __fmaf_rn(hx, A.w, __fmaf_rn(hy, Q.x, Q.y));
    e = __fsub_rn(a, fabsf(c));
    t1 = __fmaf_rn(erel, a, eabs);
}

// Exact re-evaluation of the tests of thread `t` at pixel `j` that fell inside the guard band.  The hot loop counted
// those as "not inlier"; here they get the reference's exact decision, added straight to the global counts.
template <int H, int THREADS>
__device__ __forceinline__ void k3_resolve_one(const pvv_dims &d, const pvv_ws &ws, const band_consts &bc, long long row,
                                               int hbase, int t, int j, int ox, int oy, float P, float4 A, float4 Q,
                                               uint32_t xy, unsigned long long *amb)
{
    const float cx = (float)(xy & 0xffffu), cy = (float)(xy >> 16);
    const float n1 = pvv_norm1(Q.z, Q.w);
#pragma unroll
    for (int k = 0; k < H; ++k) {
        const int h = hbase + k * THREADS + t;
        if (h >= d.hn) continue;
        const float2 hv = ws.hyp[row + h];
        float hx, hy, eabs, e, t1;
        k3_local_hyp(hv.x, hv.y, ox, oy, P, bc, hx, hy, eabs);
        k3_band(hx, hy, eabs, bc.erel, A, Q, e, t1);
        if (!(fabsf(e) > t1)) {
            if (pvv_exact_test(Q.z, Q.w, n1, cx, cy, hv.x, hv.y, d.thresh)) atomicAdd(&ws.counts[row + h], 1);
            ++*amb;
        }
    }
}


template <int H, int THREADS, bool EXACT>
__global__ void __launch_bounds__(THREADS) k_count(pvv_dims d, pvv_ws ws, band_consts bc)
{
    // per pixel two float4: {n^x, n^y, -s, tau*n^y} and {-tau*n^x, -tau*s2, raw dir.x, raw dir.y}
    __shared__ float4 sP[2 * PVV_CHUNK_PX];
    __shared__ uint32_t sXY[PVV_CHUNK_PX];
    __shared__ uint32_t s_q[EXACT ? 1 : K3_QCAP];
    __shared__ int s_item, s_P, s_qn;

    const int tid = threadIdx.x;
    const int total_items = ws.chunk_start[d.b] * d.vn;
    unsigned long long amb_local = 0;

    while (true) {
        __syncthreads();  // previous item's smem fully consumed
        if (tid == 0) { s_item = (int)atomicAdd(&ws.counters[PVV_CTR_WORK], 1u); s_P = 0; }
        __syncthreads();
        const int item = s_item;
        if (item >= total_items) break;
        const int ci = item / d.vn, vi = item - ci * d.vn;
        int lo = 0, hi = d.b;
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (ws.chunk_start[mid] <= ci) lo = mid; else hi = mid;
        }
        const int img = lo;
        const int tn = ws.tn[img];
        const long long off = ws.img_off[img];
        const int p0 = (ci - ws.chunk_start[img]) * PVV_CHUNK_PX;
        const int n = min(PVV_CHUNK_PX, tn - p0);
        const uint32_t xy0 = ws.pix_xy[off + p0];
        const int ox = (int)(xy0 & 0xffffu), oy = (int)(xy0 >> 16);

        // ---- stage the chunk: raw values + derived constants in a chunk-local frame ----
        for (int j = tid; j < n; j += THREADS) {
            const uint32_t xy = ws.pix_xy[off + p0 + j];
            const float2 dir = ws.pix_dir[(long long)vi * ws.cap + off + p0 + j];
            sXY[j] = xy;
            if (EXACT) {
                sP[2 * j + 1] = make_float4(0.f, 0.f, dir.x, dir.y);
            } else {
                const float n1 = pvv_norm1(dir.x, dir.y);
                float4 A;
                float2 B;
                if (!((double)n1 >= 1e-6) || !(n1 < 3.0e38f)) {
                    // reference skips the pixel (||n|| < 1e-6) or its cosine degenerates: never an inlier
                    A = make_float4(0.f, 0.f, -1e30f, 0.f);
                    B = make_float2(0.f, 0.f);
                } else {
                    const double inv = rsqrt((double)dir.x * dir.x + (double)dir.y * dir.y);
                    const int px = (int)(xy & 0xffffu) - ox, py = (int)(xy >> 16) - oy;
                    A.x = (float)(dir.x * inv);
                    A.y = (float)(dir.y * inv);
                    A.w = (float)((double)bc.tau * (dir.y * inv));
                    B.x = (float)(-(double)bc.tau * (dir.x * inv));
                    A.z = (float)(-((double)px * A.x + (double)py * A.y));
                    B.y = (float)(-((double)px * A.w + (double)py * B.x));
                    atomicMax(&s_P, abs(px) + abs(py));
                }
                sP[2 * j] = A;
                sP[2 * j + 1] = make_float4(B.x, B.y, dir.x, dir.y);
            }
        }
        if (!EXACT) {  // pad to a whole group with never-inlier pixels
            const int npad = (n + K3_GROUP - 1) / K3_GROUP * K3_GROUP;
            for (int j = n + tid; j < npad; j += THREADS) {
                sP[2 * j] = make_float4(0.f, 0.f, -1e30f, 0.f);
                sP[2 * j + 1] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
        __syncthreads();
        const float P = (float)s_P;
        const long long row = ((long long)img * d.vn + vi) * ws.hn_pad;

        for (int hbase = 0; hbase < d.hn; hbase += THREADS * H) {
            const bool active = hbase + (tid & ~31) < d.hn;  // warp-uniform: does this warp own hypotheses in the pass
            if (tid == 0) s_qn = 0;
            __syncthreads();
            if (active) {
                float hxr[H], hyr[H];
                int cnt[H];
#pragma unroll
                for (int k = 0; k < H; ++k) {
                    const int h = hbase + k * THREADS + tid;
                    const float2 hv = h < d.hn ? ws.hyp[row + h] : make_float2(0.f, 0.f);
                    hxr[k] = hv.x; hyr[k] = hv.y; cnt[k] = 0;
                }
                if (EXACT) {
                    for (int j = 0; j < n; ++j) {
                        const uint32_t xy = sXY[j];
                        const float4 q = sP[2 * j + 1];
                        const float cx = (float)(xy & 0xffffu), cy = (float)(xy >> 16);
                        const float n1 = pvv_norm1(q.z, q.w);
#pragma unroll
                        for (int k = 0; k < H; ++k) cnt[k] += pvv_exact_test(q.z, q.w, n1, cx, cy, hxr[k], hyr[k], d.thresh);
                    }
                } else {
                    float hx[H], hy[H], eabs[H];
#pragma unroll
                    for (int k = 0; k < H; ++k) k3_local_hyp(hxr[k], hyr[k], ox, oy, P, bc, hx[k], hy[k], eabs[k]);
                    // groups of K3_GROUP pixels share one band check / one (rare) branch
                    const float4 *pp = sP;
                    const float4 *const pend = sP + 2 * n;
                    for (; pp < pend; pp += 2 * K3_GROUP) {
                        bool clear = true;
#pragma unroll
                        for (int u = 0; u < K3_GROUP; ++u) {
                            const float4 A = pp[2 * u];
                            const float4 Q = pp[2 * u + 1];
#pragma unroll
                            for (int k = 0; k < H; ++k) {
                                float e, t1;
                                k3_band(hx[k], hy[k], eabs[k], bc.erel, A, Q, e, t1);
                                const float g = __fsub_rn(t1, e);             // < 0  <=>  e > t1: definitely an inlier
                                cnt[k] += (int)(__float_as_uint(g) >> 31);    // NaN cannot reach here (k3_local_hyp)
                                clear = clear && (fabsf(e) > t1);
                            }
                        }
                        if (!clear) {
                            // some test of this (thread, pixel group) fell inside the guard band: defer the group to the
                            // dense exact pass below (the hot loop stays divergence-free)
                            const int grp = (int)(pp - sP) / (2 * K3_GROUP);
                            const int q = atomicAdd(&s_qn, 1);
                            if (q < K3_QCAP) {
                                s_q[q] = ((uint32_t)grp << 16) | (uint32_t)tid;
                            } else {
                                for (int j = grp * K3_GROUP; j < min(n, (grp + 1) * K3_GROUP); ++j)
                                    k3_resolve_one<H, THREADS>(d, ws, bc, row, hbase, tid, j, ox, oy, P, sP[2 * j], sP[2 * j + 1],
                                                               sXY[j], &amb_local);
                            }
                        }
                    }
                }
#pragma unroll
                for (int k = 0; k < H; ++k) {
                    const int h = hbase + k * THREADS + tid;
                    if (h < d.hn && cnt[k]) atomicAdd(&ws.counts[row + h], cnt[k]);
                }
            }
            __syncthreads();
            if (!EXACT) {
                const int qn = min(s_qn, K3_QCAP);
                for (int e = tid; e < qn; e += THREADS) {
                    const uint32_t ent = s_q[e];
                    const int grp = (int)(ent >> 16), t = (int)(ent & 0xffffu);
                    for (int j = grp * K3_GROUP; j < min(n, (grp + 1) * K3_GROUP); ++j)
                        k3_resolve_one<H, THREADS>(d, ws, bc, row, hbase, t, j, ox, oy, P, sP[2 * j], sP[2 * j + 1], sXY[j],
                                                   &amb_local);
                }
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
    bc->tau = (float)tau;
    bc->erel = (float)(1.10 * etau / (1.0 - etau) + 8.0 * u);
    bc->kabs = (float)(8.0 * u * (1.0 + tau) * (1.0 + etau));
    bc->abs_floor = (float)(4e-6 * (1.0 + tau));
    return true;
}

template <int H, int THREADS, bool EXACT>
static cudaError_t launch_count_t(const pvv_dims &d, const pvv_ws &ws, const band_consts &bc, int sms, cudaStream_t st)
{
    int per_sm = 8;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_count<H, THREADS, EXACT>, THREADS, 0);
    if (per_sm < 1) per_sm = 1;
    k_count<H, THREADS, EXACT><<<sms * per_sm, THREADS, 0, st>>>(d, ws, bc);
    return cudaGetLastError();
}

cudaError_t pvv_launch_count(const pvv_dims &d, const pvv_ws &ws, int force_exact, cudaStream_t st, int *launches)
{
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    band_consts bc = {0.f, 0.f, 0.f, 0.f};
    const bool fast = !force_exact && pvv_band_for(d.thresh, &bc);
    const char *venv = getenv("PVV_K3_VARIANT");  // tuning knob for experiments
    const int variant = venv ? atoi(venv) : 0;
    cudaError_t e;
    if (!fast) e = launch_count_t<4, 128, true>(d, ws, bc, sms, st);
    else if (variant == 1) e = launch_count_t<8, 64, false>(d, ws, bc, sms, st);
    else if (variant == 2) e = launch_count_t<8, 128, false>(d, ws, bc, sms, st);
    else if (variant == 3) e = launch_count_t<2, 256, false>(d, ws, bc, sms, st);
    else e = launch_count_t<4, 128, false>(d, ws, bc, sms, st);
    ++*launches;
    return e;
}
