// pvv_vote.cu -- K2 (hypothesis generation) and K3 (fused inlier counting, the FP32-bound hot loop).
//
// K2 replaces generate_hypothesis (ransac_voting_kernel.cu:11-86) for the whole batch.
// K3 replaces  zeros([hn,vn,tn]) + voting_for_hypothesis (ransac_voting_kernel.cu:88-167) + torch.sum(inlier, 2)
//              (ransac_voting_gpu.py:563-567 / :371-374): the [hn,vn,tn] uint8 matrix is never materialised.
//
// K3 design (DESIGN.md "K3"):
//   * persistent CTAs pull (256-pixel chunk, keypoint) items from a dynamic work queue (chunks are cut in 2 or 4 when the
//     job is small); an item is decoded with one load from the chunk table K1 wrote; the raw pixels of the next item are
//     prefetched with cp.async into a second shared-memory buffer while the current item is voted on;
//   * per pixel, 6 derived floats turn the reference's cosine test into two half-plane tests through the pixel
//     (cone a > tau|c|, tau = t/sqrt(1-t^2)):  e+- = hx*p + hy*q + r,  m = min(e+, e-),  inlier <=> m > 0.
//     Pixels are stored as PAIRS so that one packed FFMA2 evaluates a term for two pixels; per test the loop issues
//     2 FFMA2 + FMNMX + 1/2 PRMT + 1/4 IADD3 (sign counting in 16-bit lanes) + 1/2 FMNMX3 (band tracker); every thread
//     keeps 4 hypotheses and their counters in registers, so the inner loop has no atomics, no shuffles, no branches and
//     no global traffic;
//   * the filter is guard-banded: a test with |m| <= T (T bounds the filter's rounding plus the uncertainty of the
//     reference's own fp32 quotient) only sets a bit in a per-thread mask; after the loop the flagged (hypothesis,
//     16-pixel group) pairs are re-evaluated by the CTA (phase A, the loop's own packed FMA chains) and the tests truly
//     inside the band get the reference's exact rounding sequence (phase B, pvv_exact_test).  Counts are bit-identical
//     to the reference kernel's (error analysis in DESIGN.md "Guard band").
#include <stdio.h>
#include <stdlib.h>

#include "pvv_internal.cuh"

namespace {

// ---- K2 -------------------------------------------------------------------------------------------
__device__ __forceinline__ void hypothesis_body(const pvv_dims &d, const int32_t *idxs, const pvv_ws &ws, float *hyp_out, int img,
                                                int e, const float *hyp_inject = nullptr)
{
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
        if (hyp_inject) {  // tests only (pvv_debug.hyp_inject): hypotheses placed by the caller
            const long long o = (((long long)img * d.hn + hi) * d.vn + vi) * 2;
            out = make_float2(hyp_inject[o], hyp_inject[o + 1]);
        } else if ((unsigned)t0 < (unsigned)tn && (unsigned)t1 < (unsigned)tn) {
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

__global__ void __launch_bounds__(256) k_hypotheses(pvv_dims d, const int32_t *idxs, pvv_ws ws, float *hyp_out,
                                                    const float *hyp_inject)
{
    pvv_pdl_enter();
    hypothesis_body(d, idxs, ws, hyp_out, blockIdx.y, blockIdx.x * blockDim.x + threadIdx.x, hyp_inject);
}

// ---- K3 -------------------------------------------------------------------------------------------
struct band_consts {
    float tau;        // t / sqrt(1 - t^2)
    float krel;       // half-width of the band per unit of max |coordinate| (covers tau's uncertainty + rounding)
    float abs_floor;  // catches ||d|| < 1e-6 (reference skips those)
    int ablate;       // experiments only (tools/k3_ablate.py builds with -DPVV_K3_ABLATE): phases to leave out
    // the two parts of krel separately (tensor-pipe kernel: they multiply different magnitudes)
    float erel;       // uncertainty of the reference's own decision, per unit of a bound on ||h - pixel||
    float kabs;       // rounding of the FP32 filter, per unit of the magnitudes in its frame
    float kabs_tc;    // rounding of the tensor-pipe filter (3xTF32 split), same unit
};
#ifndef K3_HOTVAR
#define K3_HOTVAR 0
#endif
#ifdef PVV_K3_ABLATE
#define K3_ABL(bit) (bc.ablate & (bit))
#else
#define K3_ABL(bit) 0
#endif

#ifndef K3_GROUP_PX
#define K3_GROUP_PX 16
#endif
#ifndef K3_SUB_LANES
#define K3_SUB_LANES 2
#endif
constexpr int K3_GROUP = K3_GROUP_PX;   // pixels per band-check group in the hot loop (eight packed pixel pairs): measured best of 8/16
constexpr int K3_QCAP = 512;   // deferred (thread, pixel group) entries per work item; overflow is resolved in line
constexpr int K3_GSTRIDE = 3 * (K3_GROUP / 2) + 1;  // 16-byte units per pixel group in shared memory: the pad spreads the
                                                    // groups over the banks for the per-entry reads of phase A
constexpr int K3_SP_FLOATS = 4 * K3_GSTRIDE * (PVV_CHUNK_PX / K3_GROUP);
constexpr int K3_SUB = K3_SUB_LANES;      // phase A: lanes per flagged (hypothesis, group) entry

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
__device__ __forceinline__ f32x2 ffma2(f32x2 a, f32x2 b, f32x2 c)
{
    f32x2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}

// Sign bits of two floats as two 16-bit lanes of one word: lane = 0xFFFF (-1 mod 2^16) when the value is negative, else 0.
// One PRMT in its sign-replicate mode (selector nibble 8|byte: the byte's msb copied to all 8 bits).  Words are summed
// as plain 32-bit integers (IADD3 takes two of them per instruction); the carries between the lanes are undone by
// sign_lanes_total as long as each lane saw fewer than 65536 negatives.
__device__ __forceinline__ int sign_lanes(float x, float y)
{
    int d;
    asm("prmt.b32 %0, %1, %2, 0xFFBB;" : "=r"(d) : "r"(__float_as_int(x)), "r"(__float_as_int(y)));
    return d;
}
__device__ __forceinline__ int sign_lanes_total(int acc)
{
    // acc = k0 * 0xFFFF + k1 * 0xFFFF0000 (mod 2^32), k0 / k1 = negatives seen by the low / high lane
    const uint32_t w = (uint32_t)acc;
    const uint32_t k0 = (0u - w) & 0xffffu;
    const uint32_t k1 = (0u - ((w - k0 * 0xffffu) >> 16)) & 0xffffu;
    return (int)(k0 + k1);
}

// Per-pixel constants of the filter, in the chunk-local frame (origin = middle pixel of the chunk).  With d = h - pixel,
// n^ the unit direction, a = d . n^ and c = d x n^, the reference's test cos > t is the cone a > tau |c|, i.e. the
// intersection of two half-planes through the pixel:
//   e+ = a - tau c = hx*px + hy*py + pz > 0      and      e- = a + tau c = hx*qx + hy*qy + qz > 0,
// so one test is two 2-FMA chains and m = min(e+, e-) = a - tau|c| carries the decision in its sign.
// Shared-memory layout: pixel PAIRS, 12 floats (3 x LDS.128) per pair, every constant as {pixel 2q, pixel 2q+1} so
// that one FFMA2 evaluates a term for two pixels:  [px px' py py'] [pz pz' qx qx'] [qy qy' qz qz'].
struct px_consts { float px, py, pz, qx, qy, qz; };

__device__ __forceinline__ px_consts k3_load_px(const float *sP, int j)
{
    const float *q = sP + (j >> 1) * 12 + (j / K3_GROUP) * 4 + (j & 1);
    px_consts c;
    c.px = q[0]; c.py = q[2]; c.pz = q[4]; c.qx = q[6]; c.qy = q[8]; c.qz = q[10];
    return c;
}
__device__ __forceinline__ void k3_store_px(float *sP, int j, const px_consts &c)
{
    float *q = sP + (j >> 1) * 12 + (j / K3_GROUP) * 4 + (j & 1);
    q[0] = c.px; q[2] = c.py; q[4] = c.pz; q[6] = c.qx; q[8] = c.qy; q[10] = c.qz;
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

// m = min(e+, e-) for one test, scalar form (the hot loop computes the same FMA chains with FFMA2: identical bits)
__device__ __forceinline__ float k3_m(float hx, float hy, const px_consts &c)
{
    const float ep = __fmaf_rn(hx, c.px, __fmaf_rn(hy, c.py, c.pz));
    const float em = __fmaf_rn(hx, c.qx, __fmaf_rn(hy, c.qy, c.qz));
    return fminf(ep, em);
}

// work item -> (image, keypoint, pixel range)
struct k3_item { int img, vi, p0, n, ox, oy; long long off; };

// `split` (1, 2 or 4) cuts every chunk into that many pixel ranges, each its own work item: small jobs (few chunks per
// CTA of the persistent grid) balance better; the per-item set-up is amortised over fewer pixels, so large jobs use 1.
__device__ __forceinline__ int k3_split_for(int chunks_times_vn, int grid_ctas)
{
    // measured (profiles/r01_sweep.md): worth it only below ~2.5 items per CTA; the per-item set-up and the latency of the
    // deferred passes eat the balance gain above that, and for the one-warp CTAs of small hn at any size
    return 2 * chunks_times_vn >= 5 * grid_ctas ? 1 : 2 * chunks_times_vn >= 3 * grid_ctas ? 2 : 4;
}

__device__ __forceinline__ k3_item k3_decode_item(const pvv_dims &d, const pvv_ws &ws, int item, int split = 1)
{
    k3_item it;
    const int sub = item % split;
    item /= split;
    const int ci = item / d.vn;
    it.vi = item - ci * d.vn;
    const int4 e = __ldg(&ws.chunk_tab[ci]);  // {image, pool index of the first pixel, pixels | flags, x|y<<16 of the middle pixel}
    const int span = PVV_CHUNK_PX / split;
    it.img = e.x;
    it.p0 = e.y + sub * span;
    it.n = max(0, min(span, (e.z & 0x7fff) - sub * span));  // 0: the chunk ends before this range (an empty item)
    // frame origin = the chunk's middle pixel: halves the extent P of the chunk around it, and with it the band
    PVV_CHECK(ci >= 0 && ci < ws.chunk_start[d.b] && e.x >= 0 && e.x < d.b && e.y >= 0 && (long long)e.y + (e.z & 0x7fff) <= ws.cap &&
              (e.z & 0x7fff) >= 1 && (e.z & 0x7fff) <= PVV_CHUNK_PX);
    it.ox = (int)((uint32_t)e.w & 0xffffu);
    it.oy = (int)((uint32_t)e.w >> 16);
    it.off = 0;   // p0 is a pool index already
    return it;
}

// exact-only kernel: every test with the reference's rounding sequence (force_exact, thresholds outside the filter range)
template <int H, int THREADS>
__global__ void __launch_bounds__(THREADS) k_count_exact(pvv_dims d, pvv_ws ws)
{
    __shared__ float2 sDir[PVV_CHUNK_PX];
    __shared__ uint32_t sXY[PVV_CHUNK_PX];
    __shared__ int s_item;
    pvv_pdl_enter();
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

// Guard-banded filter kernel.  The hot loop counts a test as inlier iff the sign bit of m is clear and flags, per
// (hypothesis slot, group of K3_GROUP pixels), whether any |m| <= T.  Flagged groups are re-evaluated after the loop in
// two dense phases: (A) recompute m for the tests of the group and queue the ones truly inside the band, (B) give those
// the reference's exact decision and correct the count by (exact - provisional).
// Software pipeline: while item i is being voted on, the raw pixels of item i+1 are already in flight to the other
// shared-memory buffer (cp.async) and the id of item i+2 is being fetched from the work queue, so neither the global
// load latency nor the atomic's round trip is exposed.
__device__ __forceinline__ void cp_async4(void *smem, const void *gmem)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem));
}
__device__ __forceinline__ void cp_async8(void *smem, const void *gmem)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// PSPLIT: warp groups that share the hypotheses of a pass and split the item's pixel groups between them.  HT = THREADS /
// PSPLIT threads own H hypotheses each; with few hypotheses (hn <= 128: HT = 32) PSPLIT = 2 gives a CTA two warps on the
// same shared-memory footprint (one-warp CTAs are limited to 11 per SM by their shared memory: 2.75 warps per scheduler).
template <int H, int THREADS, int MINB, int PSPLIT = 1>
__global__ void __launch_bounds__(THREADS, MINB) k_count(pvv_dims d, pvv_ws ws, band_consts bc)
{
    constexpr int HT = THREADS / PSPLIT;
    static_assert(THREADS <= 256 && H <= 8 && PVV_CHUNK_PX / K3_GROUP <= 32, "entry packing");
    static_assert(HT % 32 == 0 && HT * PSPLIT == THREADS, "a warp belongs to one pixel range");
    __shared__ __align__(16) float sP[K3_SP_FLOATS];
    __shared__ __align__(8) float2 sDir[2][PVV_CHUNK_PX];  // raw pixels, double-buffered (current / prefetched item)
    __shared__ uint32_t sXY[2][PVV_CHUNK_PX];
    __shared__ float4 sH[H * HT];  // this pass's hypotheses in the chunk-local frame: {hx, hy, T, -}
    __shared__ uint32_t s_q[K3_QCAP];   // phase A: (group << 16) | (k << 8) | thread
    __shared__ uint32_t s_q2[K3_QCAP];  // phase B: (pixel << 16) | (k << 8) | thread
    __shared__ int s_id[2], s_P[2], s_qn, s_q2n;

    pvv_pdl_enter();
    const int tid = threadIdx.x;
    const int ht = tid % HT, pg = tid / HT;  // hypothesis thread, pixel range (warp-uniform)
    const int split = THREADS >= 128 ? k3_split_for(ws.chunk_start[d.b] * d.vn, (int)gridDim.x) : 1;
    const int total_items = ws.chunk_start[d.b] * d.vn * split;
    unsigned long long amb_local = 0;
    const px_consts never = {0.f, 0.f, -1e30f, 0.f, 0.f, -1e30f};  // m = -1e30: never an inlier, never in the band

    auto prefetch = [&](const k3_item &it, int buf) {
        for (int j = tid; j < it.n; j += THREADS) {
            cp_async4(&sXY[buf][j], ws.pix_xy + it.off + it.p0 + j);
            cp_async8(&sDir[buf][j], ws.pix_dir + (long long)it.vi * ws.cap + it.off + it.p0 + j);
        }
        cp_async_commit();
    };

#ifdef PVV_K3_ABLATE
    {   // experiment: stagger the CTAs of one SM by (ablate >> 8) x 100 ns each, to break their lock-step
        const unsigned long long ns = 100ull * (unsigned)(bc.ablate >> 8) * (blockIdx.x / 148);
        unsigned long long t0, t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        do { __nanosleep(100); asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1)); } while (t1 - t0 < ns);
    }
#endif
    if (tid == 0) {
        s_id[0] = (int)atomicAdd(&ws.counters[PVV_CTR_WORK], 1u);
        s_P[0] = 0; s_P[1] = 0; s_qn = 0; s_q2n = 0;
    }
    __syncthreads();
    int item = s_id[0];
    if (item >= total_items) return;
    k3_item it = k3_decode_item(d, ws, item, split);
    prefetch(it, 0);
    if (tid == 0) s_id[1] = (int)atomicAdd(&ws.counters[PVV_CTR_WORK], 1u);

    for (int buf = 0;; buf ^= 1) {
        cp_async_wait_all();
        __syncthreads();  // raw pixels of `item` landed; s_id[buf ^ 1] visible; previous item's smem fully consumed
        const int next = s_id[buf ^ 1];
        const int n = it.n, ox = it.ox, oy = it.oy;
        const float2 *rDir = sDir[buf];
        const uint32_t *rXY = sXY[buf];

        // ---- derived constants in the chunk-local frame ----
        const int npad = (n + K3_GROUP - 1) / K3_GROUP * K3_GROUP;
        int ext = 0;  // this thread's largest |offset| from the frame origin
        for (int j = tid; j < npad; j += THREADS) {
            px_consts c = never;
            if (j < n && !K3_ABL(4)) {
                const uint32_t xy = rXY[j];
                const float2 dir = rDir[j];
                const float n1 = pvv_norm1(dir.x, dir.y);
                // the reference skips ||n|| < 1e-6; a non-finite norm degenerates its cosine: never an inlier either
                if ((double)n1 >= 1e-6 && n1 < 3.0e38f) {
                    const double inv = rsqrt((double)dir.x * dir.x + (double)dir.y * dir.y);
                    const double nx = dir.x * inv, ny = dir.y * inv, tau = (double)bc.tau;
                    const int ix = (int)(xy & 0xffffu) - ox, iy = (int)(xy >> 16) - oy;
                    c.px = (float)(nx - tau * ny);
                    c.py = (float)(ny + tau * nx);
                    c.qx = (float)(nx + tau * ny);
                    c.qy = (float)(ny - tau * nx);
                    c.pz = (float)(-((double)ix * c.px + (double)iy * c.py));
                    c.qz = (float)(-((double)ix * c.qx + (double)iy * c.qy));
                    ext = max(ext, abs(ix) + abs(iy));
                }
            }
            k3_store_px(sP, j, c);
        }
        ext = __reduce_max_sync(0xffffffffu, ext);
        if ((tid & 31) == 0 && ext) atomicMax(&s_P[buf], ext);
        // ---- prefetch the next item's raw pixels and the id of the one after it ----
        k3_item nit = it;
        if (next < total_items) {
            nit = k3_decode_item(d, ws, next, split);
            prefetch(nit, buf ^ 1);
        }
        if (tid == 0) {
            s_id[buf] = next < total_items ? (int)atomicAdd(&ws.counters[PVV_CTR_WORK], 1u) : next;
            s_P[buf ^ 1] = 0;
            s_q2n = 0;  // all threads passed phase B of the previous item (barrier above)
        }
        __syncthreads();  // constants and s_P complete
        const float P = (float)s_P[buf];
        const long long row = ((long long)it.img * d.vn + it.vi) * ws.hn_pad;
        const int ngroups = K3_ABL(2) ? 0 : npad / K3_GROUP;

        // exact decision for one in-band test, correcting the provisional count (sign bit of m clear = counted)
        auto resolve = [&](int hbase, int t, int k, int j, float m) {
            const int h = hbase + k * HT + t;
            PVV_CHECK(h >= 0 && h < d.hn && j >= 0 && j < n && k < H && t < HT);
            const float2 hv = ws.hyp[row + h];
            const float2 dir = rDir[j];
            const uint32_t xy = rXY[j];
            const int exact = pvv_exact_test(dir.x, dir.y, pvv_norm1(dir.x, dir.y), (float)(xy & 0xffffu),
                                             (float)(xy >> 16), hv.x, hv.y, d.thresh) ? 1 : 0;
            const int prov = (int)(__float_as_uint(m) >> 31) ^ 1;
            if (exact != prov) atomicAdd(&ws.counts[row + h], exact - prov);
            ++amb_local;
        };

        // pixel groups of this warp's range
        const int gper = (ngroups + PSPLIT - 1) / PSPLIT;
        const int g0 = min(ngroups, pg * gper), g1 = min(ngroups, g0 + gper);
        for (int hbase = 0; hbase < d.hn; hbase += HT * H) {
            const bool active = hbase + (ht & ~31) < d.hn;  // warp-uniform: does this warp own hypotheses in the pass
            if (hbase > 0) {
                __syncthreads();  // the previous pass's deferred entries (s_q2, sH) are consumed
                if (tid == 0) s_q2n = 0;
            }
            if (active) {
                float hx[H], hy[H], T[H];
                int neg[H];
                uint32_t dmask[H];
#pragma unroll
                for (int k = 0; k < H; ++k) {
                    const int h = hbase + k * HT + ht;
                    const float2 hv = h < d.hn ? ws.hyp[row + h] : make_float2(0.f, 0.f);
                    k3_local_hyp(hv.x, hv.y, ox, oy, P, bc, hx[k], hy[k], T[k]);
                    neg[k] = 0;
                    dmask[k] = 0u;
                    if (PSPLIT == 1 || pg == 0) sH[k * HT + ht] = make_float4(hx[k], hy[k], T[k], 0.f);  // for the deferred pass
                }
                // The hot loop has no branch and no divergence: a (slot, group) with any test inside the guard band
                // only sets a bit in a per-thread mask.
                uint32_t bit = 1u << g0;
                const ulonglong2 *pp = reinterpret_cast<const ulonglong2 *>(sP) + g0 * K3_GSTRIDE;
                for (int g = g0; g < g1; ++g, pp += K3_GSTRIDE, bit <<= 1) {
                    float mabs[H];
#pragma unroll
                    for (int k = 0; k < H; ++k) mabs[k] = __int_as_float(0x7f800000);
#pragma unroll
                    for (int u = 0; u < K3_GROUP / 2; u += 2) {  // two pixel pairs per step
                        const ulonglong2 a0 = pp[3 * u], a1 = pp[3 * u + 1], a2 = pp[3 * u + 2];
                        const ulonglong2 b0 = pp[3 * u + 3], b1 = pp[3 * u + 4], b2 = pp[3 * u + 5];
#pragma unroll
                        for (int k = 0; k < H; ++k) {
                            const f32x2 hxx = pk(hx[k], hx[k]), hyy = pk(hy[k], hy[k]);  // FFMA2 broadcast operands
                            float ep0, ep1, em0, em1, ep2, ep3, em2, em3;
                            unpk(ffma2(hxx, a0.x, ffma2(hyy, a0.y, a1.x)), ep0, ep1);
                            unpk(ffma2(hxx, a1.y, ffma2(hyy, a2.x, a2.y)), em0, em1);
                            unpk(ffma2(hxx, b0.x, ffma2(hyy, b0.y, b1.x)), ep2, ep3);
                            unpk(ffma2(hxx, b1.y, ffma2(hyy, b2.x, b2.y)), em2, em3);
                            const float m0 = fminf(ep0, em0), m1 = fminf(ep1, em1);
                            const float m2 = fminf(ep2, em2), m3 = fminf(ep3, em3);
                            // sign bits of four tests in 3 instructions (2 PRMT + IADD3) instead of 4 LEA.HI: see sign_lanes
#if K3_HOTVAR == 0
                            neg[k] += sign_lanes(m0, m1) + sign_lanes(m2, m3);
                            mabs[k] = fminf(fminf(mabs[k], fabsf(m0)), fabsf(m1));
                            mabs[k] = fminf(fminf(mabs[k], fabsf(m2)), fabsf(m3));
#else  // experiments (tools/k3_ablate.py): marginal cost of the loop's parts; counts are wrong
                            if (K3_HOTVAR == 1) neg[k] += (int)(__float_as_uint(m0) >> 31) + (int)(__float_as_uint(m1) >> 31) +
                                                          (int)(__float_as_uint(m2) >> 31) + (int)(__float_as_uint(m3) >> 31);
                            if (K3_HOTVAR == 2 || K3_HOTVAR == 5) neg[k] += sign_lanes(m0, m1) + sign_lanes(m2, m3);
                            if (K3_HOTVAR == 3 || K3_HOTVAR == 4) neg[k] ^= __float_as_int(m0) ^ __float_as_int(m1) ^ __float_as_int(m2) ^ __float_as_int(m3);
                            if (K3_HOTVAR == 1 || K3_HOTVAR == 3) {
                                mabs[k] = fminf(fminf(mabs[k], fabsf(m0)), fabsf(m1));
                                mabs[k] = fminf(fminf(mabs[k], fabsf(m2)), fabsf(m3));
                            }
                            if (K3_HOTVAR == 5) mabs[k] = fminf(mabs[k], fminf(fminf(fabsf(m0), fabsf(m1)), fminf(fabsf(m2), fabsf(m3))));
#endif
                        }
                    }
#pragma unroll
                    for (int k = 0; k < H; ++k)
                        if (!(mabs[k] > T[k])) dmask[k] |= bit;
                }
#pragma unroll
                for (int k = 0; k < H; ++k) {
                    const int h = hbase + k * HT + ht;
                    const int cnt = (g1 - g0) * K3_GROUP - sign_lanes_total(neg[k]);  // padded and never-inlier pixels have m < 0
                    if (h < d.hn && cnt && !K3_ABL(8)) atomicAdd(&ws.counts[row + h], cnt);
                }
                // queue this thread's flagged (slot, group) pairs for the dense passes below: one shared-memory atomic per
                // warp (lane counts -> warp prefix sum -> one reservation), then every lane writes its own entries
                uint32_t dm[H];
                int mine = 0;
#pragma unroll
                for (int k = 0; k < H; ++k) {
                    dm[k] = hbase + k * HT + ht < d.hn && !K3_ABL(1) ? dmask[k] : 0u;
                    mine += __popc(dm[k]);
                }
                if (__any_sync(0xffffffffu, mine)) {
                    int incl = mine;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        const int up = __shfl_up_sync(0xffffffffu, incl, o);
                        if ((tid & 31) >= o) incl += up;
                    }
                    int base = 0;
                    if ((tid & 31) == 31) base = atomicAdd(&s_qn, incl);  // lane 31 holds the warp total
                    int q = __shfl_sync(0xffffffffu, base, 31) + incl - mine;
#pragma unroll
                    for (int k = 0; k < H; ++k) {
                        uint32_t m_k = dm[k];
                        while (m_k) {
                            const int grp = __ffs(m_k) - 1;
                            m_k &= m_k - 1;
                            if (q < K3_QCAP) {
                                PVV_CHECK(q >= 0 && grp < PVV_CHUNK_PX / K3_GROUP);
                                s_q[q] = ((uint32_t)grp << 16) | ((uint32_t)k << 8) | (uint32_t)ht;
                            } else {  // queue full (pathological inputs): resolve in line
                                const float4 hl = make_float4(hx[k], hy[k], T[k], 0.f);
                                for (int j = grp * K3_GROUP; j < min(n, (grp + 1) * K3_GROUP); ++j) {
                                    const float m = k3_m(hl.x, hl.y, k3_load_px(sP, j));
                                    if (!(fabsf(m) > hl.z)) resolve(hbase, ht, k, j, m);
                                }
                            }
                            ++q;
                        }
                    }
                }
            }
            __syncthreads();
            // phase A: K3_SUB lanes per flagged (hypothesis, group) entry re-evaluate its pixels with the hot loop's own
            // packed FMA chains (bit-identical m) and queue the tests truly inside the band
            const int qtot = min(s_qn, K3_QCAP) * K3_SUB;
            for (int idx = tid; idx < qtot; idx += THREADS) {
                const uint32_t ent = s_q[idx / K3_SUB];
                const int t = (int)(ent & 0xffu), k = (int)((ent >> 8) & 0xffu), grp = (int)(ent >> 16);
                constexpr int PAIRS = K3_GROUP / 2 / K3_SUB;
                const int u0 = (idx % K3_SUB) * PAIRS;
                const float4 hl = sH[k * HT + t];
                const f32x2 hxx = pk(hl.x, hl.x), hyy = pk(hl.y, hl.y);
                const ulonglong2 *pg = reinterpret_cast<const ulonglong2 *>(sP) + grp * K3_GSTRIDE + 3 * u0;
                uint32_t inband = 0u;
#pragma unroll
                for (int u = 0; u < PAIRS; ++u) {
                    const ulonglong2 v0 = pg[3 * u], v1 = pg[3 * u + 1], v2 = pg[3 * u + 2];
                    float ep0, ep1, em0, em1;
                    unpk(ffma2(hxx, v0.x, ffma2(hyy, v0.y, v1.x)), ep0, ep1);
                    unpk(ffma2(hxx, v1.y, ffma2(hyy, v2.x, v2.y)), em0, em1);
                    if (!(fabsf(fminf(ep0, em0)) > hl.z)) inband |= 1u << (2 * u);
                    if (!(fabsf(fminf(ep1, em1)) > hl.z)) inband |= 2u << (2 * u);
                }
                while (inband) {
                    const int j = grp * K3_GROUP + 2 * u0 + __ffs(inband) - 1;
                    inband &= inband - 1;
                    if (j >= n) continue;  // padded pixels (m = -1e30) only get here when T is infinite
                    const int q = atomicAdd(&s_q2n, 1);
                    if (q < K3_QCAP) s_q2[q] = ((uint32_t)j << 16) | ((uint32_t)k << 8) | (uint32_t)t;
                    else resolve(hbase, t, k, j, k3_m(hl.x, hl.y, k3_load_px(sP, j)));
                }
            }
            __syncthreads();
            if (tid == 0) s_qn = 0;  // every thread has read it; the next pushes come after another barrier
            // phase B: exact decisions for the tests inside the band
            const int q2 = min(s_q2n, K3_QCAP);
            for (int idx = tid; idx < q2; idx += THREADS) {
                const uint32_t ent = s_q2[idx];
                const int t = (int)(ent & 0xffu), k = (int)((ent >> 8) & 0xffu), j = (int)(ent >> 16);
                const float4 hl = sH[k * HT + t];
                resolve(hbase, t, k, j, k3_m(hl.x, hl.y, k3_load_px(sP, j)));
            }
        }
        if (next >= total_items) break;
        item = next;
        it = nit;
    }
    if (amb_local) atomicAdd((unsigned long long *)&ws.counters[PVV_CTR_AMB_LO], amb_local);
}

#include "pvv_vote_tc.cuh"
#include "pvv_vote_tc2.cuh"
}  // namespace

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
    // tensor-pipe filter (pvv_vote_tc.cuh): 3xTF32 split products.  Per product the dropped lo*lo term (4u) and the two
    // split tails (2u each), the split tail of r (2u), an 8-term accumulation taken as one truncation (2u) per term on
    // the running maximum (16u), plus the 4.5u of input / constant rounding shared with the FP32 filter: <= 31u(1+tau)M.
    // Measured on the B200 (tools/tc_probe.cu): 5.5u against the fp32 constants.  Used: 40u.
    const double kabs_tc = 40.0 * u * (1.0 + tau) * (1.0 + etau);
    bc->erel = (float)(erel * (1.0 + 4.0 * u));
    bc->kabs = (float)(kabs * (1.0 + 4.0 * u));
    bc->kabs_tc = (float)(kabs_tc * (1.0 + 4.0 * u));
    bc->abs_floor = (float)(4e-6 * (1.0 + tau));
    bc->ablate = 0;
#ifdef PVV_K3_ABLATE
    if (const char *e = getenv("PVV_K3_ABLATE")) bc->ablate = atoi(e);
#endif
    return true;
}

template <int H, int THREADS, int MINB, int PSPLIT = 1>
static cudaError_t launch_count_t(const pvv_dims &d, const pvv_ws &ws, const band_consts &bc, int sms, cudaStream_t st)
{
    int per_sm = 8;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_count<H, THREADS, MINB, PSPLIT>, THREADS, 0);
    if (per_sm < 1) per_sm = 1;
    pvv_launch_pdl(k_count<H, THREADS, MINB, PSPLIT>, dim3(sms * per_sm), dim3(THREADS), st, d, ws, bc);
    return cudaGetLastError();
}

// Which kernel counts (measured, profiles/r02_k3_kernel_crossover_by_hn.log, tools/crossover_small.py): above 128 hypotheses
// the tensor-pipe kernel (pvv_vote_tc.cuh: tcgen05 tf32 MMA for the two linear forms, CUDA cores for the sign count and the
// band) -- since its producer warps work as two groups side by side it wins from 160 hypotheses on at every batch size (1 - 32
// images: 25 against 44 us, 189 against 214 us) and is 15 - 25 % faster at 512; at 128 the FP32-pipe kernel with one-warp CTAs
// (every thread owns 4 hypotheses) is level with it on large batches (142 against 153 us with the pre-pass) and stays for
// those; small jobs of any hypothesis count go to the tensor-pipe kernel (see pvv_count_variant).
// pvv_debug.k3_variant overrides the choice (n > 0 forces variant n - 1; 0..6 and 11..13: FP32-pipe shapes, 7 / 8: tensor
// pipe, 9: tensor pipe with 3 producer warps, 10: the experimental queue-and-defer pair): a tuning and test knob
// (tools/k3_variants.py, tests/test_gpu_parity.py).  -1: the exact-only kernel.
static int pvv_count_variant(const pvv_dims &d, const pvv_ws &ws, int force_exact, int k3_variant, band_consts *bc)
{
    if (force_exact || !pvv_band_for(d.thresh, bc)) return -1;
    // Up to 128 hypotheses the choice goes by job size.  The host only knows an upper bound of the pooled pixels (the pool's
    // capacity), hence of the units of work: where even that bound is small the tensor-pipe kernel wins (b = 1: 24 against
    // 35 us for K2 + K3; the reference's shipped setting, 128 hypotheses and max_num = 100, at b = 32: 20 against 25 us); on
    // large batches with full pools the one-warp FP32-pipe CTAs do (b = 32: 142 against 153 us).  tools/crossover_tiny.py,
    // tools/crossover_small.py; the crossing is near 16 images of configs[1] = a bound of ~17 000 units.
    const long long units_bound = (ws.cap / PVV_CHUNK_PX + d.b) * d.vn;
    int v = k3_variant > 0 ? k3_variant - 1 : (d.hn <= 128 && units_bound > 9000 ? 2 : 7);
    // the second-generation tensor-pipe kernel indexes hyp / counts with 32 bits in its queue entries
    if (v == 10 && (long long)d.b * d.vn * ((d.hn + 3) / 4 * 4) >= (1LL << 31)) v = 7;
    return v;
}

// K2.  When the tensor-pipe kernel will count, its per-pixel constants pre-pass rides in the same launch (independent
// work on the same input: the two overlap instead of queueing).
cudaError_t pvv_launch_hypotheses(const pvv_dims &d, const int32_t *idxs, const pvv_ws &ws, float *hyp_out, int force_exact,
                                  int k3_variant, const float *hyp_inject, cudaStream_t st, int *launches)
{
    band_consts bc = {};
    const int hyp_bx = (d.vn * d.hn + 255) / 256;
    const int cv = pvv_count_variant(d, ws, force_exact, k3_variant, &bc);
    if (cv >= 7 && cv <= 10) {
        int dev = 0, sms = 148;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        const long long pblocks = (ws.cap + 255) / 256;  // grid-stride: the pixel count is only known on the device
        const int consts_bx = (int)(pblocks < 2 * sms ? pblocks : 2 * sms);
        pvv_launch_pdl(k_hypotheses_and_consts, dim3((unsigned)(hyp_bx * d.b + consts_bx * d.vn)), dim3(256), st, d, idxs, ws,
                       hyp_out, hyp_inject, bc.tau, hyp_bx, consts_bx);
    } else {
        pvv_launch_pdl(k_hypotheses, dim3(hyp_bx, d.b), dim3(256), st, d, idxs, ws, hyp_out, hyp_inject);
    }
    ++*launches;
    return cudaGetLastError();
}

// tensor-pipe kernel: one 768-thread CTA per SM (all 512 TMEM columns), dynamic shared memory, chained by PDL like the
// rest; its constants pre-pass ran with K2
template <int PW>
static cudaError_t launch_count_tc(const pvv_dims &d, const pvv_ws &ws, const band_consts &bc, int sms, cudaStream_t st)
{
    cudaError_t e = cudaFuncSetAttribute(k_count_tc<PW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(tc_smem));
    if (e != cudaSuccess) return e;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(sms);
    cfg.blockDim = dim3(tc_cfg<PW>::threads);
    cfg.dynamicSmemBytes = sizeof(tc_smem);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, k_count_tc<PW>, d, ws, bc);
}

// second-generation tensor-pipe kernel + its deferred-pass kernel (pvv_vote_tc2.cuh)
template <int NB>
static cudaError_t launch_count_tc2(const pvv_dims &d, const pvv_ws &ws, const band_consts &bc, int sms, cudaStream_t st, int *launches)
{
    cudaError_t e = cudaFuncSetAttribute(k_count_tc2<NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(t2_smem<NB>));
    if (e != cudaSuccess) return e;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(sms < ws.tcq_slices ? sms : ws.tcq_slices);
    cfg.blockDim = dim3(t2_cfg<NB>::threads);
    cfg.dynamicSmemBytes = sizeof(t2_smem<NB>);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if ((e = cudaLaunchKernelEx(&cfg, k_count_tc2<NB>, d, ws, bc)) != cudaSuccess) return e;
    const int parts = 16;
    pvv_launch_pdl(k_count_tc_deferred, dim3(ws.tcq_slices * parts), dim3(T2_DEF_THREADS), st, d, ws, bc, parts);
    ++*launches;
    return cudaGetLastError();
}

cudaError_t pvv_launch_count(const pvv_dims &d, const pvv_ws &ws, int force_exact, int k3_variant, cudaStream_t st,
                             int *launches)
{
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    band_consts bc = {};
    const int variant = pvv_count_variant(d, ws, force_exact, k3_variant, &bc);
    const bool fast = variant >= 0;
    cudaError_t e;
    if (!fast) {
        int per_sm = 8;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_count_exact<4, 128>, 128, 0);
        pvv_launch_pdl(k_count_exact<4, 128>, dim3(sms * (per_sm < 1 ? 1 : per_sm)), dim3(128), st, d, ws);
        e = cudaGetLastError();
    }
    else if (variant == 7 || variant == 8 || variant == 9) {
        // 7 producer warps (variant 9 forces 3: until the seventh warp's lane took over the staging of the next unit's inputs
        // they were 4 % faster where the deferred passes are rare, thresh <= 0.995; now 7 win there too: 293 against 299 us)
        const bool few = variant == 9;
        e = few ? launch_count_tc<3>(d, ws, bc, sms, st) : launch_count_tc<7>(d, ws, bc, sms, st);
    }
    else if (variant == 10) e = launch_count_tc2<4>(d, ws, bc, sms, st, launches);
    else if (variant == 11) e = launch_count_t<4, 64, 10, 2>(d, ws, bc, sms, st);
    else if (variant == 12) e = launch_count_t<4, 128, 5, 4>(d, ws, bc, sms, st);
    else if (variant == 13) e = launch_count_t<4, 128, 5, 2>(d, ws, bc, sms, st);
    else if (variant == 1) e = launch_count_t<4, 64, 10>(d, ws, bc, sms, st);
    else if (variant == 2) e = launch_count_t<4, 32, 20>(d, ws, bc, sms, st);
    else if (variant == 3) e = launch_count_t<4, 128, 6>(d, ws, bc, sms, st);
    else if (variant == 6) e = launch_count_t<4, 128, 4>(d, ws, bc, sms, st);
    else if (variant == 4) e = launch_count_t<4, 128, 7>(d, ws, bc, sms, st);
    else if (variant == 5) e = launch_count_t<4, 128, 8>(d, ws, bc, sms, st);
    else e = launch_count_t<4, 128, 5>(d, ws, bc, sms, st);  // 96 registers, no spills, 5 CTAs per SM: measured best (6 CTAs spill)
    ++*launches;
    return e;
}
