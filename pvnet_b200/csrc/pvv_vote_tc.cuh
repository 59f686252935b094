// pvv_vote_tc.cuh -- K3, tensor-pipe variant (included by pvv_vote.cu inside its anonymous namespace).
//
// The filter's two half-plane forms  e+- = hx*p + hy*q + r  are a rank-3 bilinear form in (hypothesis, pixel): per work
// item an E[hypotheses x 2*pixels] matrix = H[hyp x K] . P[2*pixels x K]^T.  The FP32 pipe spent 2 FFMA2 per test on it;
// here the 5th-generation tensor cores produce it with ONE tcgen05.mma (kind::tf32, M = 128 hypotheses, N = 128 columns =
// 64 pixels x {e+, e-}, K = 8) per tile, accumulators in TMEM, and the CUDA cores keep only what is not a contraction:
// m = min(e+, e-), sign counting and the guard-band tracker (2.5 instead of 4.9 issued instructions per test).
//
// Precision: TF32 has an 11-bit significand, so every fp32 factor is split x = x_hi + x_lo (both exactly representable
// in TF32) and the K = 8 slots carry the three significant cross terms of each product plus the split constant term:
//      A row (hypothesis) = [hx_hi hx_hi hx_lo hy_hi | hy_hi hy_lo 1    1   ]
//      B row (pixel, +-)  = [p_hi  p_lo  p_hi  q_hi  | q_lo  q_hi  r_hi r_lo]
// TF32 x TF32 products are exact in the fp32 accumulator's 24 bits; what is lost is x_lo*y_lo (2^-22 relative), the
// tail of the split (2^-23) and the accumulation rounding of the 8-term sum.  tools/tc_probe.cu measures the total on
// the B200 in units of u(1+tau)M (u = 2^-24, M = |hx|+|hy|+P): the guard band of this kernel is built on a bound with
// margin over that measurement (pvv_band_for: kabs_tc), and is only 10-20 % wider than the FP32 filter's at t = 0.999
// because the band is dominated by the uncertainty of the REFERENCE's own fp32 quotient, not by the filter's rounding.
// The tensor-pipe result is never trusted inside the band: a flagged (hypothesis, 16-pixel group) is left out of the
// provisional count altogether and re-decided by the FP32 filter (phase A, proved bound) and, inside ITS band, by the
// reference's exact rounding sequence (phase B).  Counts stay bit-identical to the reference kernel's.
//
// CTA = 8 voting warps + 1 MMA warp.  The voting warps stage the operands in shared memory (canonical K-major,
// non-swizzled core matrices: 8 rows x 16 B; LBO = byte stride between the two K halves, SBO = between 8-row groups:
// verified on the B200 by tools/tc_probe.cu), one elected thread of the MMA warp issues the tcgen05.mma of tile i+1 into
// the other TMEM buffer while the voting warps read tile i with tcgen05.ld (32x32b: lane = hypothesis row) -- the
// tensor pipe (~70 cycles per tile) hides completely behind the ~300 cycles of CUDA-core work per tile.
#pragma once

#ifdef PVV_K3_ABLATE
#define K3_ABLV bc.ablate
#else
#define K3_ABLV 0
#endif

constexpr int TC_VOTERS = 256;              // voting threads (8 warps)
constexpr int TC_THREADS = TC_VOTERS + 32;  // + the MMA warp
constexpr int TC_PASS_H = 512;              // hypotheses per pass (4 M-blocks of 128)
constexpr int TC_TILE_PX = 64;              // pixels per MMA tile (N = 128 columns)
constexpr int TC_TMEM_COLS = 256;           // two accumulator buffers of 128 columns
constexpr long long TC_SPIN_CYCLES = 4000000000ll;  // ~2 s

struct tc_smem {
    uint4 A[2][TC_PASS_H];            // hypotheses, [K half][row]: 16 KB
    uint4 B[2][2 * PVV_CHUNK_PX];     // pixels x {+,-}, [K half][row = 2*pixel + plane]: 16 KB
    float4 H[TC_PASS_H];              // local hypotheses {hx, hy, T of the FP32 filter, T of the tensor-pipe filter}
    float P[K3_SP_FLOATS];            // FP32 constants in the FP32 kernel's pair layout (phase A)
    float2 dir[2][PVV_CHUNK_PX];      // raw pixels, double-buffered
    uint32_t xy[2][PVV_CHUNK_PX];
    uint32_t q[K3_QCAP];              // phase A entries: (group << 16) | hypothesis within the pass
    uint32_t q2[K3_QCAP];             // phase B entries: (pixel << 16) | hypothesis within the pass
    unsigned long long bar_full[2], bar_empty[2], bar_ab;
    uint32_t tmem;
    int ctl_mb, ctl_nb, ctl_stop;
    int id[2], ext[2], qn, q2n;
};

__device__ __forceinline__ uint32_t tc_saddr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void tc_bar_init(unsigned long long *b, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(tc_saddr(b)), "r"(count));
}
__device__ __forceinline__ void tc_bar_arrive(unsigned long long *b)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tc_saddr(b)) : "memory");
}
// Bounded wait: in a correct run the limit (seconds) is never reached; a broken pipeline traps instead of hanging the GPU.
__device__ __forceinline__ uint32_t tc_bar_try(uint32_t bar_saddr, uint32_t parity)
{
    uint32_t done;
    asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.b32 %0, 1, 0, P1;\n\t}"
                 : "=r"(done) : "r"(bar_saddr), "r"(parity) : "memory");
    return done;
}
__device__ __noinline__ void tc_bar_wait_slow(uint32_t bar_saddr, uint32_t parity)
{
    const long long t0 = clock64();
    while (!tc_bar_try(bar_saddr, parity))
        if (clock64() - t0 > TC_SPIN_CYCLES) __trap();
}
__device__ __forceinline__ void tc_bar_wait(uint32_t bar_saddr, uint32_t parity)
{
    if (!tc_bar_try(bar_saddr, parity)) tc_bar_wait_slow(bar_saddr, parity);
}
__device__ __forceinline__ void tc_bar_wait(unsigned long long *b, uint32_t parity) { tc_bar_wait(tc_saddr(b), parity); }
__device__ __forceinline__ void tc_bar_arrive(uint32_t bar_saddr)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar_saddr) : "memory");
}
__device__ __forceinline__ void tc_voters_sync() { asm volatile("bar.sync 1, %0;" ::"n"(TC_VOTERS) : "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }

__device__ __forceinline__ uint64_t tc_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes)
{
    // K-major, SWIZZLE_NONE, descriptor version 1 (sm_100): start >> 4 | LBO >> 4 << 16 | SBO >> 4 << 32 | 1 << 46
    return (uint64_t)((saddr >> 4) & 0x3fffu) | ((uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16) |
           ((uint64_t)((sbo_bytes >> 4) & 0x3fffu) << 32) | (1ull << 46);
}

// x = hi + lo (+ a remainder below 2^-22 |x|), both exactly representable in TF32 (low 13 bits zero)
__device__ __forceinline__ void tc_split(float x, uint32_t &hi, uint32_t &lo)
{
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hi) : "f"(x));
    const float rest = __fsub_rn(x, __uint_as_float(hi));  // exact
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(lo) : "f"(rest));
}

__device__ __forceinline__ void tc_ld32(uint32_t taddr, uint32_t (&v)[32])
{
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
          "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
          "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr));
}

// One group of 16 pixels of one hypothesis: v[2i] = e+, v[2i+1] = e- of pixel i.  Adds the group's negatives (sign
// lanes) to `neg` unless any |m| <= T, in which case only `bit` is set in `dmask`.
__device__ __forceinline__ void tc_group(const uint32_t (&v)[32], float T, int &neg, uint32_t &dmask, uint32_t bit)
{
    float m[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) m[i] = fminf(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]));
    int g = 0;
    float mabs = __int_as_float(0x7f800000);
#pragma unroll
    for (int i = 0; i < 16; i += 2) {
        g += sign_lanes(m[i], m[i + 1]);
        mabs = fminf(fminf(mabs, fabsf(m[i])), fabsf(m[i + 1]));
    }
    const bool flagged = !(mabs > T);
    neg += flagged ? 0 : g;
    dmask |= flagged ? bit : 0u;
}

// One 128-column tile of this thread's hypothesis row: wait for the accumulator, pull this warp's 64 columns into
// registers, hand the TMEM buffer back, then vote on the two 16-pixel groups.
__device__ __forceinline__ void tc_tile(uint32_t a_full, uint32_t a_empty, uint32_t tmem_row, uint32_t &t, int lane, float T,
                                        int &neg, uint32_t &dm, uint32_t bit, int abl = 0)
{
    const uint32_t tb = t & 1u;
    if (!(abl & 32)) tc_bar_wait(a_full + 8u * tb, (t >> 1) & 1u);
    ++t;
    tc_fence_after();
    uint32_t v0[32], v1[32];
    tc_ld32(tmem_row + tb * 128u, v0);
    tc_ld32(tmem_row + tb * 128u + 32u, v1);
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");  // warp-collective: every lane's columns have landed
    tc_fence_before();
    if (lane == 0 && !(abl & 32)) tc_bar_arrive(a_empty + 8u * tb);  // the accumulator is in registers: buffer free for the next MMA
    if (abl & 16) { neg += v0[0] ^ v1[0]; return; }
    tc_group(v0, T, neg, dm, bit);
    tc_group(v1, T, neg, dm, bit << 1);
}

__global__ void __launch_bounds__(TC_THREADS, 2) k_count_tc(pvv_dims d, pvv_ws ws, band_consts bc, float krel_tc)
{
    extern __shared__ __align__(128) unsigned char tc_smem_raw[];
    tc_smem &S = *reinterpret_cast<tc_smem *>(tc_smem_raw);
    static_assert(K3_GROUP == 16 && PVV_CHUNK_PX == 256 && TC_TILE_PX == 64 && TC_VOTERS == PVV_CHUNK_PX, "tile / group geometry");

    pvv_pdl_enter();
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int total_items = ws.chunk_start[d.b] * d.vn;

    if (tid == 0) {
        tc_bar_init(&S.bar_full[0], 1); tc_bar_init(&S.bar_full[1], 1);
        tc_bar_init(&S.bar_empty[0], TC_VOTERS / 32); tc_bar_init(&S.bar_empty[1], TC_VOTERS / 32);
        tc_bar_init(&S.bar_ab, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        S.id[0] = (int)atomicAdd(&ws.counters[PVV_CTR_WORK], 1u);
        S.ext[0] = 0; S.ext[1] = 0; S.qn = 0; S.q2n = 0; S.ctl_stop = 0;
    }
    if (warp == TC_VOTERS / 32) {  // the MMA warp owns the TMEM allocation
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tc_saddr(&S.tmem)), "r"(TC_TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = S.tmem;

    if (warp == TC_VOTERS / 32) {
        // ================= MMA issuer: one thread =================
        if (lane == 0) {
            const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(128 >> 3) << 17) | ((128u >> 4) << 24);
            uint32_t t = 0, ab_parity = 0;
            while (true) {
                tc_bar_wait(&S.bar_ab, ab_parity);
                ab_parity ^= 1u;
                if (*(volatile int *)&S.ctl_stop) break;
                const int mbn = *(volatile int *)&S.ctl_mb, nbn = *(volatile int *)&S.ctl_nb;
                tc_fence_after();
                for (int mb = 0; mb < mbn && !K3_ABL(32); ++mb) {
                    const uint64_t da = tc_desc(tc_saddr(&S.A[0][mb * 128]), sizeof(S.A[0]), 128);
                    for (int nb = 0; nb < nbn; ++nb, ++t) {
                        const uint32_t buf = t & 1u;
                        if (t >= 2) tc_bar_wait(&S.bar_empty[buf], ((t >> 1) - 1u) & 1u);
                        tc_fence_after();
                        const uint64_t db = tc_desc(tc_saddr(&S.B[0][nb * 128]), sizeof(S.B[0]), 128);
                        asm volatile(
                            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                            "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem + buf * 128u),
                            "l"(da), "l"(db), "r"(idesc), "r"(0));
                        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                                         tc_saddr(&S.bar_full[buf])) : "memory");
                    }
                }
            }
        }
        __syncwarp();
    } else {
        // ================= voting warps =================
        const int q4 = warp & 3;     // TMEM lane quarter = rows 32*q4 .. +31 of the M-block
        const int chalf = warp >> 2; // column half of the tile: pixels 32*chalf .. +31 of its 64
        unsigned long long amb_local = 0;
        const px_consts never = {0.f, 0.f, -1e30f, 0.f, 0.f, -1e30f};
        uint32_t t = 0;  // tile counter, in lock-step with the MMA issuer's
        const uint32_t a_full = tc_saddr(&S.bar_full[0]), a_empty = tc_saddr(&S.bar_empty[0]);
        const uint32_t tmem_row = tmem + ((uint32_t)(q4 * 32) << 16) + (uint32_t)(chalf * 64);

        auto prefetch = [&](const k3_item &it, int buf) {
            if (tid < it.n) {  // it.n <= PVV_CHUNK_PX = TC_VOTERS
                cp_async4(&S.xy[buf][tid], ws.pix_xy + it.off + it.p0 + tid);
                cp_async8(&S.dir[buf][tid], ws.pix_dir + (long long)it.vi * ws.cap + it.off + it.p0 + tid);
            }
            cp_async_commit();
        };

        int item = S.id[0];
        if (item < total_items) {
            k3_item it = k3_decode_item(d, ws, item);
            prefetch(it, 0);
            if (tid == 0) S.id[1] = (int)atomicAdd(&ws.counters[PVV_CTR_WORK], 1u);

            for (int buf = 0;; buf ^= 1) {
                cp_async_wait_all();
                tc_voters_sync();  // raw pixels landed; S.id[buf ^ 1] visible; the previous item's shared memory is consumed
                const int next = S.id[buf ^ 1];
                const int n = it.n, ox = it.ox, oy = it.oy;
                const float2 *rDir = S.dir[buf];
                const uint32_t *rXY = S.xy[buf];
                const int nbn = (n + TC_TILE_PX - 1) / TC_TILE_PX;

                // ---- pixel constants: FP32 (phase A) and their TF32 split as B rows ----
                int ext = 0;
                if (tid < nbn * TC_TILE_PX) {
                    const int j = tid;
                    px_consts c = never;
                    if (j < n && !K3_ABL(4)) {
                        const uint32_t xy = rXY[j];
                        const float2 dir = rDir[j];
                        const float n1 = pvv_norm1(dir.x, dir.y);
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
                            ext = abs(ix) + abs(iy);
                        }
                    }
                    k3_store_px(S.P, j, c);
                    uint32_t ph, pl, qh, ql, rh, rl;
                    tc_split(c.px, ph, pl); tc_split(c.py, qh, ql); tc_split(c.pz, rh, rl);
                    S.B[0][2 * j] = make_uint4(ph, pl, ph, qh);
                    S.B[1][2 * j] = make_uint4(ql, qh, rh, rl);
                    tc_split(c.qx, ph, pl); tc_split(c.qy, qh, ql); tc_split(c.qz, rh, rl);
                    S.B[0][2 * j + 1] = make_uint4(ph, pl, ph, qh);
                    S.B[1][2 * j + 1] = make_uint4(ql, qh, rh, rl);
                }
                ext = __reduce_max_sync(0xffffffffu, ext);
                if (lane == 0 && ext) atomicMax(&S.ext[buf], ext);
                // ---- prefetch the next item's raw pixels and the id of the one after it ----
                k3_item nit = it;
                if (next < total_items) {
                    nit = k3_decode_item(d, ws, next);
                    prefetch(nit, buf ^ 1);
                }
                if (tid == 0) {
                    S.id[buf] = next < total_items ? (int)atomicAdd(&ws.counters[PVV_CTR_WORK], 1u) : next;
                    S.ext[buf ^ 1] = 0;
                }
                tc_voters_sync();  // constants and S.ext complete
                const float P = (float)S.ext[buf];
                const long long row = ((long long)it.img * d.vn + it.vi) * ws.hn_pad;

                auto resolve = [&](int h, int j) {  // the reference's own decision for one test
                    const float2 hv = ws.hyp[row + h];
                    const float2 dir = rDir[j];
                    const uint32_t xy = rXY[j];
                    if (pvv_exact_test(dir.x, dir.y, pvv_norm1(dir.x, dir.y), (float)(xy & 0xffffu), (float)(xy >> 16), hv.x,
                                       hv.y, d.thresh))
                        atomicAdd(&ws.counts[row + h], 1);
                    ++amb_local;
                };
                // phase A body for one (hypothesis, group): FP32 filter on its 16 pixels; in-band tests go to phase B
                auto refilter = [&](int hbase, int hl, int grp, int u0, int pairs) {
                    const float4 hh = S.H[hl];
                    const f32x2 hxx = pk(hh.x, hh.x), hyy = pk(hh.y, hh.y);
                    const ulonglong2 *pg = reinterpret_cast<const ulonglong2 *>(S.P) + grp * K3_GSTRIDE + 3 * u0;
                    int sure = 0;
                    uint32_t inband = 0u;
                    for (int u = 0; u < pairs; ++u) {
                        const ulonglong2 v0 = pg[3 * u], v1 = pg[3 * u + 1], v2 = pg[3 * u + 2];
                        float ep0, ep1, em0, em1;
                        unpk(ffma2(hxx, v0.x, ffma2(hyy, v0.y, v1.x)), ep0, ep1);
                        unpk(ffma2(hxx, v1.y, ffma2(hyy, v2.x, v2.y)), em0, em1);
                        const float m0 = fminf(ep0, em0), m1 = fminf(ep1, em1);
                        if (!(fabsf(m0) > hh.z)) inband |= 1u << (2 * u); else sure += m0 > 0.f;
                        if (!(fabsf(m1) > hh.z)) inband |= 2u << (2 * u); else sure += m1 > 0.f;
                    }
                    if (sure) atomicAdd(&ws.counts[row + hbase + hl], sure);
                    while (inband) {
                        const int j = grp * K3_GROUP + 2 * u0 + __ffs(inband) - 1;
                        inband &= inband - 1;
                        if (j >= n) continue;  // padded pixels (m = -1e30) only get here when T is infinite
                        const int q = atomicAdd(&S.q2n, 1);
                        if (q < K3_QCAP) S.q2[q] = ((uint32_t)j << 16) | (uint32_t)hl;
                        else resolve(hbase + hl, j);
                    }
                };

                for (int hbase = 0; hbase < d.hn; hbase += TC_PASS_H) {
                    const int mbn = min(TC_PASS_H / 128, (d.hn - hbase + 127) / 128);
                    if (hbase > 0) tc_voters_sync();  // the previous pass's deferred entries (S.q2, S.H) are consumed
                    if (tid == 0) S.q2n = 0;
                    // ---- hypotheses of this pass in the chunk-local frame: A rows and band half-widths ----
                    for (int hl = tid; hl < mbn * 128 && !K3_ABL(64); hl += TC_VOTERS) {
                        const int h = hbase + hl;
                        const float2 hv = h < d.hn ? ws.hyp[row + h] : make_float2(0.f, 0.f);
                        float hx = __fsub_rn(hv.x, (float)ox), hy = __fsub_rn(hv.y, (float)oy), Tf, Tt;
                        const float mm = __fadd_rn(__fadd_rn(fabsf(hx), fabsf(hy)), P);
                        if (mm < 1e15f) {
                            Tf = __fmaf_rn(bc.krel, mm, bc.abs_floor);
                            Tt = __fmaf_rn(krel_tc, mm, bc.abs_floor);
                        } else {  // far-away / non-finite hypothesis: every test takes the exact path
                            hx = 0.f; hy = 0.f; Tf = Tt = __int_as_float(0x7f800000);
                        }
                        S.H[hl] = make_float4(hx, hy, Tf, Tt);
                        uint32_t xh, xl, yh, yl;
                        tc_split(hx, xh, xl); tc_split(hy, yh, yl);
                        S.A[0][hl] = make_uint4(xh, xh, xl, yh);
                        S.A[1][hl] = make_uint4(yh, yl, 0x3f800000u, 0x3f800000u);
                    }
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // operands visible to the tensor pipe
                    tc_voters_sync();
                    if (tid == 0) {
                        S.ctl_mb = mbn; S.ctl_nb = nbn;
                        __threadfence_block();
                        tc_bar_arrive(&S.bar_ab);
                    }
                    // ---- hot loop: tiles out of TMEM ----
                    uint32_t dmask = 0u;  // bit mb*8 + nb*2 + g: flagged 16-pixel group of this thread's hypothesis in M-block mb
                    for (int mb = 0; mb < mbn; ++mb) {
                        const int hl = mb * 128 + q4 * 32 + lane;
                        const float T = S.H[hl].w;
                        int neg = 0;
                        uint32_t dm = 0u;
                        if (nbn == PVV_CHUNK_PX / TC_TILE_PX) {  // full chunk (the common case): four tiles, unrolled
                            tc_tile(a_full, a_empty, tmem_row, t, lane, T, neg, dm, 1u, K3_ABLV);
                            tc_tile(a_full, a_empty, tmem_row, t, lane, T, neg, dm, 4u, K3_ABLV);
                            tc_tile(a_full, a_empty, tmem_row, t, lane, T, neg, dm, 16u, K3_ABLV);
                            tc_tile(a_full, a_empty, tmem_row, t, lane, T, neg, dm, 64u, K3_ABLV);
                        } else {
                            for (int nb = 0; nb < nbn; ++nb) tc_tile(a_full, a_empty, tmem_row, t, lane, T, neg, dm, 1u << (2 * nb), K3_ABLV);
                        }
                        dmask |= dm << (mb * 8);
                        const int cnt = K3_GROUP * (2 * nbn - __popc(dm)) - sign_lanes_total(neg);
                        if (hbase + hl < d.hn && cnt) atomicAdd(&ws.counts[row + hbase + hl], cnt);
                    }
                    // ---- queue the flagged (hypothesis, group) pairs: one shared-memory reservation per warp ----
                    {
                        uint32_t dm = K3_ABL(1) ? 0u : dmask;
                        // hypotheses beyond hn own no counts
#pragma unroll
                        for (int mb = 0; mb < TC_PASS_H / 128; ++mb)
                            if (hbase + mb * 128 + q4 * 32 + lane >= d.hn) dm &= ~(0xffu << (mb * 8));
                        const int mine = __popc(dm);
                        if (__any_sync(0xffffffffu, mine)) {
                            int incl = mine;
#pragma unroll
                            for (int o = 1; o < 32; o <<= 1) {
                                const int up = __shfl_up_sync(0xffffffffu, incl, o);
                                if (lane >= o) incl += up;
                            }
                            int base = 0;
                            if (lane == 31) base = atomicAdd(&S.qn, incl);
                            int q = __shfl_sync(0xffffffffu, base, 31) + incl - mine;
                            while (dm) {
                                const int bitn = __ffs(dm) - 1;
                                dm &= dm - 1;
                                const int hl = (bitn >> 3) * 128 + q4 * 32 + lane;
                                const int grp = ((bitn >> 1) & 3) * 4 + chalf * 2 + (bitn & 1);  // 16-pixel group within the chunk
                                if (q < K3_QCAP) S.q[q] = ((uint32_t)grp << 16) | (uint32_t)hl;
                                else refilter(hbase, hl, grp, 0, K3_GROUP / 2);  // queue full (pathological inputs): in line
                                ++q;
                            }
                        }
                    }
                    tc_voters_sync();
                    // phase A: K3_SUB lanes per flagged entry
                    const int qtot = min(S.qn, K3_QCAP) * K3_SUB;
                    for (int idx = tid; idx < qtot; idx += TC_VOTERS) {
                        const uint32_t ent = S.q[idx / K3_SUB];
                        constexpr int PAIRS = K3_GROUP / 2 / K3_SUB;
                        refilter(hbase, (int)(ent & 0xffffu), (int)(ent >> 16), (idx % K3_SUB) * PAIRS, PAIRS);
                    }
                    tc_voters_sync();
                    if (tid == 0) S.qn = 0;
                    // phase B: exact decisions for the tests inside the FP32 filter's band
                    const int q2 = min(S.q2n, K3_QCAP);
                    for (int idx = tid; idx < q2; idx += TC_VOTERS) {
                        const uint32_t ent = S.q2[idx];
                        resolve(hbase + (int)(ent & 0xffffu), (int)(ent >> 16));
                    }
                }
                if (next >= total_items) break;
                item = next;
                it = nit;
            }
        }
        if (amb_local) atomicAdd((unsigned long long *)&ws.counters[PVV_CTR_AMB_LO], amb_local);
        tc_voters_sync();
        if (tid == 0) {
            S.ctl_stop = 1;
            __threadfence_block();
            tc_bar_arrive(&S.bar_ab);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == TC_VOTERS / 32)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TC_TMEM_COLS));
}
