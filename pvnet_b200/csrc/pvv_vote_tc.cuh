// pvv_vote_tc.cuh -- K3, tensor-pipe kernel (included by pvv_vote.cu inside its anonymous namespace; shipped for hn > 256).
//
// The filter's decision quantity  m = a - tau|c|  (a = d . n^, c = d x n^, d = hypothesis - pixel; m = min(e+, e-) of the
// FP32-pipe kernel's two half-plane forms) is built from two forms that are LINEAR in the hypothesis and in per-pixel
// constants:  a = hx*nx + hy*ny + ra,  tau c = hx*(tau ny) + hy*(-tau nx) + rc  -- a rank-3 bilinear form in (hypothesis,
// pixel): per unit of work an E[hypotheses x 2*pixels] matrix = H[hyp x K] . P[2*pixels x K]^T.  The FP32-pipe kernel
// (k_count) spends 2 FFMA2 per test on it; here the 5th-generation tensor cores produce it with ONE tcgen05.mma
// (kind::tf32, M = 128 hypotheses, N = 256 columns = 128 pixels x {a, tau c}, K = 8) per tile, accumulators in TMEM, and
// the CUDA cores keep only what is not a contraction: m = a - |tau c| (one FADD), sign counting, the guard-band tracker.
//
// Precision: TF32 has an 11-bit significand, so every fp32 factor is split x = x_hi + x_lo (both exactly representable
// in TF32) and the K = 8 slots carry the three significant cross terms of each product plus the split constant term:
//      A row (hypothesis) = [hx_hi hx_hi hx_lo hy_hi | hy_hi hy_lo 1    1   ]
//      B row (pixel, form) = [p_hi  p_lo  p_hi  q_hi  | q_lo  q_hi  r_hi r_lo]      (p, q, r) = (nx, ny, ra) or (tau ny, -tau nx, rc)
// TF32 x TF32 products are exact in the fp32 accumulator's 24 bits; what is lost is x_lo*y_lo (2^-22 relative), the
// tail of the split (2^-23) and the accumulation rounding of the 8-term sum.  tools/tc_probe.cu measures the total on
// the B200 in units of u(1+tau)M (u = 2^-24, M = |hx|+|hy|+P): 5.5; the guard band of this kernel is built on 40
// (pvv_band_for: kabs_tc), and is only ~15 % wider than the FP32 filter's at t = 0.999 because the band is dominated
// by the uncertainty of the REFERENCE's own fp32 quotient, not by the filter's rounding.
// The tensor-pipe result is never trusted inside the band: a flagged (hypothesis, 16-pixel group) is left out of the
// provisional count altogether and re-decided by the FP32 filter (phase A, proved bound) and, inside ITS band, by the
// reference's exact rounding sequence (phase B).  Counts stay bit-identical to the reference kernel's.
//
// One 768-thread CTA per SM (all 512 TMEM columns: two 256-column accumulator buffers), three roles, setmaxnreg:
//   * 16 voting warps (96 registers): warp = (TMEM lane quarter, column quarter); per 16-pixel group one tcgen05.ld
//     32x32b.x32 (lane = hypothesis row), software-pipelined so the load of group g+1 is in flight while group g is voted
//     on; counts accumulate in shared memory per ROW (image, keypoint, pass) and reach global memory once per row;
//   * 1 MMA warp: one elected thread issues the tcgen05.mma of tile i+2 as soon as the voters have pulled tile i into
//     registers (mbarriers full / empty; tcgen05.commit); operands are canonical K-major, non-swizzled core matrices
//     (8 rows x 16 B; LBO = byte stride between the two K halves, SBO = between 8-row groups: verified by tc_probe.cu);
//   * 7 producer warps (48 registers): stage the raw inputs of unit k+1 with 1-D bulk copies (TMA), build the B rows of unit k
//     (pixel constants from the pre-pass that rides in K2's launch, translated into the row's frame and split), build the A rows
//     once per row, and run the deferred passes (phase A / B) of unit k-2 -- the voters never leave their loop.
// Units are dealt as contiguous ranges in (keypoint, pass, chunk) order, so a CTA changes row ~3 times per launch.
#pragma once

#ifdef PVV_K3_ABLATE
#define K3_ABLV bc.ablate
#else
#define K3_ABLV 0
#endif

constexpr int TC_VOTER_WARPS = 16;                 // voting warps: 4 TMEM lane quarters x 4 column quarters of a tile
constexpr int TC_VOTERS = TC_VOTER_WARPS * 32;
constexpr int TC_MMA_WARP = TC_VOTER_WARPS;        // one elected thread issues every tcgen05.mma
// Producer warps (operands of the next unit, deferred passes of the previous) are a template parameter PW of the kernel:
//   PW = 7: 768 threads, launch 80 registers -> setmaxnreg voters 96 / others 48 (512 x 96 + 256 x 48 = 768 x 80);
//   PW = 3: 640 threads, launch 96 registers -> voters 104 / others 64 (512 x 104 + 128 x 64 = 640 x 96).
// The sums must be equal: setmaxnreg.inc waits until the CTA's pool has the registers.  Seven warps keep up with the
// deferred passes at t = 0.999 (4.3e-4 of the tests go through them); where those are rare (t <= 0.995) three suffice and
// the voters' 8 extra registers save them re-materialised addresses: 4 % faster (profiles/r01_k3_tc_ablation.log).
#ifndef TC_RV7
#define TC_RV7 96
#define TC_RO7 48
#endif
template <int PW> struct tc_cfg {
    static constexpr int prods = PW * 32;
    static constexpr int threads = TC_VOTERS + 32 + PW * 32;
    static constexpr int regs_voter = PW == 7 ? TC_RV7 : 104, regs_other = PW == 7 ? TC_RO7 : 64;
    static_assert(PW == 7 || PW == 3, "register split worked out for 7 and 3 producer warps");
};
constexpr int TC_STAGES = 3;                       // unit k lives in stage k % 3: produced | voted on | deferred passes
constexpr int TC_PASS_H = 512;                     // hypotheses per unit (4 M-blocks of 128)
constexpr int TC_TILE_PX = 128;                    // pixels per MMA tile (N = 256 columns)
constexpr int TC_TMEM_COLS = 512;                  // two accumulator buffers of 256 columns: the whole TMEM of the SM
constexpr uint32_t TC_SPIN_ROUNDS = 1u << 22;       // x 64 polls: seconds of continuous polling by one warp
#ifndef TC_SLEEP_MMA
#define TC_SLEEP_MMA 0
#endif
#ifndef TC_SLEEP_VOTER
#define TC_SLEEP_VOTER 32
#endif
#ifndef TC_SLEEP_PROD
#define TC_SLEEP_PROD 400
#endif

// One unit of work = (256-pixel chunk, keypoint, pass of <= 512 hypotheses).  Units are ordered (keypoint, pass, chunk)
// and every CTA takes one contiguous range of them, so consecutive units of a CTA share their ROW = (image, keypoint,
// pass): the hypothesis operand (A rows, in a frame whose origin is the middle pixel of the row's first chunk in this
// CTA) is built once per row, not once per unit.
constexpr int TC_ROWS = 4;    // row buffers (a row is rewritten 3 rows later: its last unit's deferred passes are over)
constexpr int TC_RAWS = 4;    // raw-input buffers (unit k+1 is being staged while unit k-2 is in its deferred passes)
struct tc_row {
    uint4 A[2][TC_PASS_H];            // hypotheses, [K half][row]: 16 KB
    float4 H[TC_PASS_H];              // {hx, hy, |hx|+|hy| (inf: every test of this hypothesis takes the exact path), -}
    float2 raw[TC_PASS_H];            // the hypotheses as K2 wrote them (phase B)
    int cnt[TC_PASS_H];               // inlier counts of the row, accumulated over this CTA's units of it: one global add per
                                      // hypothesis and ROW instead of one per hypothesis, unit and voting warp
};
struct tc_raw {                       // raw inputs of one unit, staged by cp.async one iteration ahead
    float4 pq[PVV_CHUNK_PX];          // frame-independent constants (pre-pass)
    float2 dir[PVV_CHUNK_PX];         // direction vectors and coordinates (phase B)
    uint32_t xy[PVV_CHUNK_PX];
};
struct tc_stage {
    uint4 B[2][2 * PVV_CHUNK_PX];     // pixels x {a, tau c}, [K half][row = 2*pixel + form]: 16 KB
    float P[K3_SP_FLOATS];            // FP32 constants in the FP32 kernel's pair layout (phase A)
    uint32_t q[K3_QCAP];              // flagged entries: (group << 16) | hypothesis within the pass
    long long cnt_row;                // offset of the (image, keypoint) row + pass in counts
    float ext;                        // largest |offset| of a pixel of the unit from the row's frame origin
    float ucx, ucy, uext;             // the unit's own centre (its middle pixel) in the row's frame, and the largest |offset|
                                      // of a pixel of the unit from it: |h - centre| + uext bounds ||h - pixel||
    int qn, stop, mbn, nbn, n, hcount, row, last;   // last: the row ends with this unit (its counts are flushed)
};
// -DPVV_TC_PROF: cycle stamps of CTA 0's first units, per role, printed when the kernel ends (development builds only)
#ifdef PVV_TC_PROF
#define TC_PROF_UNITS 48
#define TC_PROF_SLOTS 20
__device__ long long g_tc_prof[TC_PROF_UNITS][TC_PROF_SLOTS];
#define TC_STAMP(k, slot) do { if (blockIdx.x == 0 && (k) < TC_PROF_UNITS) g_tc_prof[(k)][(slot)] = clock64(); } while (0)
#else
#define TC_STAMP(k, slot) ((void)0)
#endif

struct tc_smem {
    tc_row rows[TC_ROWS];
    tc_stage st[TC_STAGES];
    tc_raw raw[TC_RAWS];
    unsigned long long bar_ab[TC_STAGES], bar_done[TC_STAGES], bar_full[2], bar_empty[2];
    unsigned long long bar_free[TC_STAGES];   // (7 producer warps) the deferred pass of the stage's unit is over: the stage can be rebuilt
    unsigned long long bar_in[TC_RAWS];   // raw inputs of unit k have landed in raw[k % TC_RAWS] (bulk copies: complete_tx)
    long long off[8];                 // ring: pool offset of the image of unit k (slot k & 7), fetched two iterations ahead
    int4 ent[8];                      // ring: chunk-table entry of unit k, fetched three iterations ahead (x = -1: no unit)
    uint32_t tmem;
    int ext, uext;
};

__device__ __forceinline__ void cp_async16(void *smem, const void *gmem)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem));
}
__device__ __forceinline__ uint32_t tc_saddr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void tc_bar_init(unsigned long long *b, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(tc_saddr(b)), "r"(count));
}
__device__ __forceinline__ void tc_bar_arrive(unsigned long long *b)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tc_saddr(b)) : "memory");
}
// Bounded wait: in a correct run the limit is never reached; a broken pipeline traps instead of hanging the GPU.
__device__ __forceinline__ uint32_t tc_bar_try(uint32_t bar_saddr, uint32_t parity)
{
    uint32_t done;
    asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.b32 %0, 1, 0, P1;\n\t}"
                 : "=r"(done) : "r"(bar_saddr), "r"(parity) : "memory");
    return done;
}
// SLEEP_NS: back-off between two polls.  A waiting warp must not eat the issue slots of the voting warps (before the
// back-off, 22 % of the kernel's executed instructions were polls): producers, who run a unit ahead, sleep long; voters
// wait for the tensor pipe only briefly; the MMA issuer, whose wake-up latency is on the path of every tile, does not sleep.
template <int SLEEP_NS>
__device__ __forceinline__ void tc_bar_wait_slow(uint32_t bar_saddr, uint32_t parity)
{
    // The bound counts THIS warp's polls, not wall-clock cycles: a context that is time-sliced out (another process on the
    // GPU, MPS, a debugger) does not advance it, so a correct run cannot trip it.  A legitimate wait lasts microseconds.
    for (uint32_t rounds = 0;; ++rounds) {
#pragma unroll 1
        for (int i = 0; i < 64; ++i) {
            if (tc_bar_try(bar_saddr, parity)) return;
            if (SLEEP_NS) __nanosleep(SLEEP_NS);
        }
#ifndef PVV_NO_WATCHDOG
        if (rounds > TC_SPIN_ROUNDS) __trap();
#endif
    }
}
template <int SLEEP_NS = TC_SLEEP_VOTER>
__device__ __forceinline__ void tc_bar_wait(uint32_t bar_saddr, uint32_t parity)
{
    if (!tc_bar_try(bar_saddr, parity)) tc_bar_wait_slow<SLEEP_NS>(bar_saddr, parity);
}
template <int SLEEP_NS = TC_SLEEP_VOTER>
__device__ __forceinline__ void tc_bar_wait(unsigned long long *b, uint32_t parity) { tc_bar_wait<SLEEP_NS>(tc_saddr(b), parity); }
__device__ __forceinline__ void tc_bar_arrive(uint32_t bar_saddr)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar_saddr) : "memory");
}
// 1-D bulk copy global -> shared (TMA), completion counted in bytes on an mbarrier of this CTA
__device__ __forceinline__ void tc_bulk_g2s(void *smem, const void *gmem, uint32_t bytes, unsigned long long *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(tc_saddr(smem)), "l"(gmem), "r"(bytes), "r"(tc_saddr(bar)) : "memory");
}
__device__ __forceinline__ void tc_expect_tx(unsigned long long *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tc_saddr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }

// one lane of a converged warp (elect.sync): the lane that issues tcgen05.mma / tcgen05.commit
__device__ __forceinline__ bool tc_elect_one()
{
    uint32_t pred;
    asm volatile("{\n\t.reg .pred P1;\n\telect.sync _|P1, 0xffffffff;\n\tselp.b32 %0, 1, 0, P1;\n\t}" : "=r"(pred));
    return pred != 0;
}

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

// One group of 16 pixels of one hypothesis: v[2i] = a, v[2i+1] = tau c of pixel i (m = a - |tau c|).  Adds the group's negatives (sign
// lanes) to `neg` unless any |m| <= T, in which case only `bit` is set in `dmask`.
__device__ __forceinline__ void tc_group(const uint32_t (&v)[32], float T, int &neg, uint32_t &dmask, uint32_t bit)
{
    float m[16];
#pragma unroll
    // m = a - |tau c| = min(e+, e-): one FADD with an |.| modifier on the FMA pipe (the ALU pipe is the one this kernel fills)
    for (int i = 0; i < 16; ++i) m[i] = __fsub_rn(__uint_as_float(v[2 * i]), fabsf(__uint_as_float(v[2 * i + 1])));
    int g = 0;
    float mabs = __int_as_float(0x7f800000);
#pragma unroll
    for (int i = 0; i < 16; i += 2) {
        g += sign_lanes(m[i], m[i + 1]);
        mabs = fminf(fminf(mabs, fabsf(m[i])), fabsf(m[i + 1]));
    }
    // clean group (every |m| > T): its negatives count; otherwise only its bit is set -- two predicated instructions
    asm("{\n\t.reg .pred p;\n\tsetp.gt.f32 p, %2, %3;\n\t@p add.s32 %0, %0, %4;\n\t@!p or.b32 %1, %1, %5;\n\t}"
        : "+r"(neg), "+r"(dmask) : "f"(mabs), "f"(T), "r"(g), "r"(bit));
}

// Pre-pass (one thread per (keypoint, pooled pixel)): the four frame-independent constants of the two linear forms
//   a = d . n^ = hx*nx + hy*ny + ra          tau c = tau (d x n^) = hx*(tau ny) + hy*(-tau nx) + rc
// (e+- = a -+ tau c, so m = min(e+, e-) = a - |tau c|), each correctly rounded from fp64: {nx, ny, tau ny, -tau nx}.  All-zero marks a pixel the
// reference never counts (||n|| < 1e-6, or a non-finite norm).
__device__ __forceinline__ void pixel_consts_body(const pvv_dims &d, const pvv_ws &ws, float tau_f, int vi, long long first,
                                                  long long stride)
{
    const long long total = ws.img_off[d.b - 1] + ws.tn[d.b - 1];
    const float2 *dirs = ws.pix_dir + (long long)vi * ws.cap;
    float4 *out = ws.pix_pq + (long long)vi * ws.cap;
    const double tau = (double)tau_f;
    for (long long i = first; i < total; i += stride) {
        const float2 dir = dirs[i];
        const float n1 = pvv_norm1(dir.x, dir.y);
        float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
        if ((double)n1 >= 1e-6 && n1 < 3.0e38f) {
            const double inv = rsqrt((double)dir.x * dir.x + (double)dir.y * dir.y);
            const double nx = dir.x * inv, ny = dir.y * inv;
            c = make_float4((float)nx, (float)ny, (float)(tau * ny), (float)(-tau * nx));
        }
        out[i] = c;
    }
}

// K2 and the pre-pass in one launch (1-D grid): blocks [0, hyp_bx * b) are K2's (hyp_bx per image), the next
// consts_bx * vn blocks stride over the pooled pixels of one keypoint plane each.
__global__ void __launch_bounds__(256) k_hypotheses_and_consts(pvv_dims d, const int32_t *idxs, pvv_ws ws, float *hyp_out,
                                                               const float *hyp_inject, float tau_f, int hyp_bx, int consts_bx)
{
    pvv_pdl_enter();
    const int bx = blockIdx.x;
    if (bx < hyp_bx * d.b) {
        hypothesis_body(d, idxs, ws, hyp_out, bx / hyp_bx, (bx % hyp_bx) * 256 + threadIdx.x, hyp_inject);
    } else {
        const int cb = bx - hyp_bx * d.b;
        pixel_consts_body(d, ws, tau_f, cb / consts_bx, (long long)(cb % consts_bx) * 256 + threadIdx.x, (long long)consts_bx * 256);
    }
}

template <int PW>
__global__ void __launch_bounds__(tc_cfg<PW>::threads, 1) k_count_tc(pvv_dims d, pvv_ws ws, band_consts bc)
{
    constexpr int TC_PROD_WARPS = PW, TC_PRODS = tc_cfg<PW>::prods, TC_THREADS = tc_cfg<PW>::threads;
    constexpr int TC_REGS_VOTER = tc_cfg<PW>::regs_voter, TC_REGS_OTHER = tc_cfg<PW>::regs_other;
    // Seven producer warps work as two groups side by side: five build the operands of unit k (four build rows, the fifth's
    // first lane stages the inputs of unit k + 1), two run the deferred pass of the unit the voters finished last.  One after
    // the other these took ~3 700 cycles per unit (profiles/r02_k_count_tc_timeline.md), which bounded the kernel below 256
    // hypotheses; side by side the longer of the two (~2 200) does.
    constexpr bool SPLIT = PW == 7;
    constexpr int TC_ROW_WARPS = 5, TC_DEF_WARPS = 2;            // SPLIT only
    constexpr int TC_ROWG = TC_ROW_WARPS * 32, TC_DEFG = TC_DEF_WARPS * 32;
    extern __shared__ __align__(128) unsigned char tc_smem_raw[];
    tc_smem &S = *reinterpret_cast<tc_smem *>(tc_smem_raw);
    static_assert(K3_GROUP == 16 && PVV_CHUNK_PX == 256 && TC_TILE_PX == 128, "tile / group geometry");

    pvv_pdl_enter();
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    if (tid == 0) {
        for (int i = 0; i < TC_STAGES; ++i) {
            tc_bar_init(&S.bar_ab[i], SPLIT ? TC_ROW_WARPS : TC_PROD_WARPS);
            tc_bar_init(&S.bar_done[i], TC_VOTER_WARPS);
            tc_bar_init(&S.bar_free[i], TC_DEF_WARPS);
        }
        for (int i = 0; i < TC_RAWS; ++i) tc_bar_init(&S.bar_in[i], 1);
        tc_bar_init(&S.bar_full[0], 1); tc_bar_init(&S.bar_full[1], 1);
        tc_bar_init(&S.bar_empty[0], TC_VOTER_WARPS); tc_bar_init(&S.bar_empty[1], TC_VOTER_WARPS);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        S.ext = 0; S.uext = 0;
    }
    for (int i = tid; i < TC_ROWS * TC_PASS_H; i += TC_THREADS) S.rows[i / TC_PASS_H].cnt[i % TC_PASS_H] = 0;
    if (warp == TC_MMA_WARP) {  // the MMA warp owns the TMEM allocation
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tc_saddr(&S.tmem)), "r"(TC_TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = S.tmem;

    if (warp < TC_VOTER_WARPS) {
        // ================= voting warps: TMEM -> registers -> counts, nothing else =================
        asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(TC_REGS_VOTER));
        const int q4 = warp & 3;   // TMEM lane quarter = rows 32*q4 .. +31 of the M-block
        const int cq = warp >> 2;  // column quarter of the tile: pixels 32*cq .. +31 of its 128
        uint32_t a_full = tc_saddr(&S.bar_full[0]), a_empty = tc_saddr(&S.bar_empty[0]);
        uint32_t tmem_row = tmem + ((uint32_t)(q4 * 32) << 16) + (uint32_t)(cq * 64);
        // opaque to the compiler: kept in registers instead of being rematerialised from tid / the shared window base
        uint32_t lane0 = lane == 0;
        asm volatile("" : "+r"(a_full), "+r"(a_empty), "+r"(tmem_row), "+r"(lane0));
        uint32_t t = 0;  // tile counter, in lock-step with the MMA issuer's
        for (uint32_t k = 0;; ++k) {
            const uint32_t s = k % TC_STAGES;
            tc_stage &T = S.st[s];
            if (tid == 0) TC_STAMP(k, 8);
            tc_bar_wait(&S.bar_ab[s], (k / TC_STAGES) & 1u);
            if (T.stop) break;
            if (tid == 0) TC_STAMP(k, 9);
            const int mbn = T.mbn, nbn = T.nbn, hcount = T.hcount;
            const float ext = T.ext, ucx = T.ucx, ucy = T.ucy, uext = T.uext;
            tc_row &R = S.rows[T.row];
            uint32_t dmask = 0u;  // bit mb*4 + nb*2 + g: flagged 16-pixel group of this thread's hypothesis in M-block mb
            const int abl = K3_ABLV;
            uint32_t v0[32], v1[32];
#ifdef PVV_TC_PROF
            long long pf_wait = 0, pf_ld = 0, pf_t;
#define TC_PF_BEGIN() (pf_t = clock64())
#define TC_PF_END(acc) (acc += clock64() - pf_t)
#else
#define TC_PF_BEGIN() ((void)0)
#define TC_PF_END(acc) ((void)0)
#endif
            auto full_wait = [&](uint32_t tt) {
                TC_PF_BEGIN();
                if (!(abl & 32)) tc_bar_wait(a_full + 8u * (tt & 1u), (tt >> 1) & 1u);
                tc_fence_after();
                TC_PF_END(pf_wait);
            };
            auto landed = [&] { TC_PF_BEGIN(); asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); TC_PF_END(pf_ld); };
            auto release = [&](uint32_t tt) {  // both groups of tile tt are in registers: its TMEM buffer is free
                tc_fence_before();
                if (lane0 && !(abl & 32)) tc_bar_arrive(a_empty + 8u * (tt & 1u));
            };
            auto vote = [&](const uint32_t (&v)[32], float Tt, int &neg, uint32_t &dm, uint32_t bit) {
                if (abl & 16) neg += v[0] ^ v[31]; else tc_group(v, Tt, neg, dm, bit);
            };
            // Both groups of a tile are in registers before its first vote, so its TMEM buffer goes back at once (the MMA of
            // tile t+2 has three votes of slack); the first group of the next tile is loaded under the vote on this tile's second
            // group, its second group right after it (one exposed tcgen05.ld per tile: measured 1 % better than pipelining both).
            if (!(abl & 2048)) {
                full_wait(t);
                tc_ld32(tmem_row + (t & 1u) * 256u, v0);
                tc_ld32(tmem_row + (t & 1u) * 256u + 32u, v1);
                landed();
                release(t);
            }
            for (int mb = 0; mb < mbn && !(abl & 2048); ++mb) {
                const int hl = mb * 128 + q4 * 32 + lane;
                const float4 hh = R.H[hl];
                const float mu = __fadd_rn(__fadd_rn(fabsf(__fsub_rn(hh.x, ucx)), fabsf(__fsub_rn(hh.y, ucy))), uext);
                const float Tt = __fmaf_rn(bc.erel, mu, __fmaf_rn(bc.kabs_tc, __fadd_rn(hh.z, ext), bc.abs_floor));
                int neg = 0;
                uint32_t dm = 0u;
                const bool more = mb + 1 < mbn;
                vote(v0, Tt, neg, dm, 1u);
                if (nbn == 2) {
                    full_wait(t + 1);
                    tc_ld32(tmem_row + ((t + 1) & 1u) * 256u, v0);
                    vote(v1, Tt, neg, dm, 2u);
                    tc_ld32(tmem_row + ((t + 1) & 1u) * 256u + 32u, v1);
                    landed();
                    release(t + 1);
                    vote(v0, Tt, neg, dm, 4u);
                    t += 2;
                    if (more) {
                        full_wait(t);
                        tc_ld32(tmem_row + (t & 1u) * 256u, v0);
                    }
                    vote(v1, Tt, neg, dm, 8u);
                    if (more) {
                        tc_ld32(tmem_row + (t & 1u) * 256u + 32u, v1);
                        landed();
                        release(t);
                    }
                } else {
                    t += 1;
                    if (more) {
                        full_wait(t);
                        tc_ld32(tmem_row + (t & 1u) * 256u, v0);
                    }
                    vote(v1, Tt, neg, dm, 2u);
                    if (more) {
                        tc_ld32(tmem_row + (t & 1u) * 256u + 32u, v1);
                        landed();
                        release(t);
                    }
                }
                const int cnt = K3_GROUP * (2 * nbn - __popc(dm)) - sign_lanes_total(neg);
                if (hl < hcount) {
                    if (cnt && !K3_ABL(8)) atomicAdd(&R.cnt[hl], cnt);
                    dmask |= dm << (mb * 4);
                }
            }
            if (tid == 0) TC_STAMP(k, 10);
#ifdef PVV_TC_PROF
            if (tid == 0 && blockIdx.x == 0 && k < TC_PROF_UNITS) { g_tc_prof[k][13] = pf_wait; g_tc_prof[k][14] = pf_ld; }
#endif
            // queue the flagged (hypothesis, group) pairs for the producers' deferred passes: one reservation per warp
            if (K3_ABL(1)) dmask = 0u;
            // (a lane with flagged groups reserves its queue slots itself: ~3.5 entries per warp and unit, so a handful of
            // shared-memory atomics cost less than the warp-wide prefix sum + one atomic they replace -- every voter warp
            // runs this tail at the same moment, so its latency is dead time for the whole SM.  Pulling the next unit's first
            // tile in front of this tail, to overlap the two, was measured slower: 324 against 308 us)
            while (dmask) {
                const int bitn = __ffs(dmask) - 1;
                dmask &= dmask - 1;
                const int hl = (bitn >> 2) * 128 + q4 * 32 + lane;
                const int grp = ((bitn >> 1) & 1) * 8 + cq * 2 + (bitn & 1);  // 16-pixel group within the chunk
                const int q = atomicAdd(&T.qn, 1);
                if (q < K3_QCAP) {
                    T.q[q] = ((uint32_t)grp << 16) | (uint32_t)hl;
                } else {  // queue full (pathological inputs): the reference's own decision for the whole group, in line
                    const tc_raw &W = S.raw[k % TC_RAWS];
                    const float2 hv = R.raw[hl];
                    int c = 0;
                    for (int j = grp * K3_GROUP; j < min(T.n, (grp + 1) * K3_GROUP); ++j) {
                        const float2 dir = W.dir[j];
                        const uint32_t xy = W.xy[j];
                        c += pvv_exact_test(dir.x, dir.y, pvv_norm1(dir.x, dir.y), (float)(xy & 0xffffu), (float)(xy >> 16),
                                            hv.x, hv.y, d.thresh) ? 1 : 0;
                    }
                    if (c) atomicAdd(&R.cnt[hl], c);
                }
            }
            __syncwarp();
            if (lane == 0) tc_bar_arrive(&S.bar_done[s]);
            if (tid == 0) TC_STAMP(k, 11);
        }
    } else {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(TC_REGS_OTHER));
        if (warp == TC_MMA_WARP) {
            // ================= MMA issuer: one warp, warp-uniform control flow, one elected lane issues =================
            // (Round 1 ran this loop in a single diverged thread: the compiler then shuttles every operand of the uniform-
            // datapath instructions through R2UR / ELECT / BRA.U.ANY sequences, ~50 dependent instructions per tile on the
            // critical path of a two-deep accumulator ring.  With warp-uniform loops the descriptors live in uniform registers.)
            {
                const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(256 >> 3) << 17) | ((128u >> 4) << 24);
                const uint32_t full0 = tc_saddr(&S.bar_full[0]), empty0 = tc_saddr(&S.bar_empty[0]);
                uint32_t t = 0;
                for (uint32_t k = 0;; ++k) {
                    const uint32_t s = k % TC_STAGES;
                    tc_stage &T = S.st[s];
                    tc_bar_wait<TC_SLEEP_MMA>(&S.bar_ab[s], (k / TC_STAGES) & 1u);
                    if (T.stop) break;
                    if (lane == 0) TC_STAMP(k, 12);
                    const int mbn = T.mbn, nbn = T.nbn;
                    const tc_row &R = S.rows[T.row];
                    tc_fence_after();
                    // descriptors advance linearly: 128 (256) rows x 16 B = 128 (256) units of 16 B in the start-address field
                    uint64_t da = tc_desc(tc_saddr(&R.A[0][0]), sizeof(R.A[0]), 128);
                    const uint64_t db0 = tc_desc(tc_saddr(&T.B[0][0]), sizeof(T.B[0]), 128);
                    for (int mb = 0; mb < mbn && !K3_ABL(32); ++mb, da += 128) {
                        uint64_t db = db0;
                        for (int nb = 0; nb < nbn; ++nb, ++t, db += 256) {
                            const uint32_t buf = t & 1u;
                            if (t >= 2) tc_bar_wait<TC_SLEEP_MMA>(empty0 + 8u * buf, ((t >> 1) - 1u) & 1u);
                            tc_fence_after();
                            if (tc_elect_one()) {
                                asm volatile(
                                    "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                    "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem + buf * 256u),
                                    "l"(da), "l"(db), "r"(idesc), "r"(0));
                                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                                                 full0 + 8u * buf) : "memory");
                            }
                            __syncwarp();
                        }
                    }
                }
            }
            __syncwarp();
        } else {
            // ================= producers: operands of unit k, then the deferred passes of unit k - 2 =================
            const int ptid = tid - (TC_VOTERS + 32);
            // barrier 1: the operand group (all producers when the roles are not split); barrier 2: the deferred-pass group
            auto pbar = [] { asm volatile("bar.sync 1, %0;" ::"n"(SPLIT ? TC_ROWG : TC_PRODS) : "memory"); };
            auto dbar = [] { asm volatile("bar.sync 2, %0;" ::"n"(TC_DEFG) : "memory"); };
            const bool def_role = SPLIT && ptid >= TC_ROWG;           // this thread belongs to the deferred-pass group
            const int gt = def_role ? ptid - TC_ROWG : ptid;          // thread index within the group that runs a deferred pass
            constexpr int GN = SPLIT ? TC_DEFG : TC_PRODS;            // ... and that group's size
            // With 7 producer warps the last one does not build operands: its first lane issues the look-ahead loads and the
            // bulk copies of the next unit's inputs (~1500 cycles of one thread: address arithmetic, expect_tx, 3 - 4 bulk
            // copies) while the other six build this unit's rows -- 256 pixels take two rounds with 192 threads as with 224.
            // (Profile of CTA 0, -DPVV_TC_PROF: the staging used to sit on the critical path of thread 0, in front of its share
            // of the rows.)  With 3 producer warps every thread is needed for the rows.
            constexpr int SPT = SPLIT ? TC_ROWG - 32 : 0;              // the staging thread
            constexpr int TC_PB = SPLIT ? TC_ROWG - 32 : TC_PRODS;     // threads that build operands: ptid < TC_PB
            unsigned long long amb_local = 0;
            const px_consts never = {0.f, 0.f, -1e30f, 0.f, 0.f, -1e30f};

            auto resolve = [&](const tc_stage &T, const tc_raw &W, int hl, int j) {  // the reference's own decision for one test
                const float2 hv = S.rows[T.row].raw[hl];
                const float2 dir = W.dir[j];
                const uint32_t xy = W.xy[j];
                if (pvv_exact_test(dir.x, dir.y, pvv_norm1(dir.x, dir.y), (float)(xy & 0xffffu), (float)(xy >> 16), hv.x, hv.y,
                                   d.thresh))
                    atomicAdd(&S.rows[T.row].cnt[hl], 1);
                ++amb_local;
            };
            // deferred passes of one unit: (A) the FP32 filter re-decides every test of a flagged (hypothesis, group) --
            // the tensor-pipe result is not used inside its band at all; (B) tests inside the FP32 filter's own band get
            // the reference's exact rounding sequence
            auto deferred = [&](uint32_t u) {
                const uint32_t s = u % TC_STAGES;
                tc_stage &T = S.st[s];
                const tc_raw &W = S.raw[u % TC_RAWS];
                tc_row &R = S.rows[T.row];
                tc_bar_wait<TC_SLEEP_PROD>(&S.bar_done[s], (u / TC_STAGES) & 1u);
                if (gt == 0) TC_STAMP(u + 2, 6);
                const int qtot = min(T.qn, K3_QCAP) * K3_SUB;
                for (int idx = gt; idx < qtot; idx += GN) {
                    const uint32_t ent = T.q[idx / K3_SUB];
                    constexpr int PAIRS = K3_GROUP / 2 / K3_SUB;
                    const int hl = (int)(ent & 0xffffu), grp = (int)(ent >> 16), u0 = (idx % K3_SUB) * PAIRS;
                    const float4 hh = R.H[hl];
                    const float mu = __fadd_rn(__fadd_rn(fabsf(__fsub_rn(hh.x, T.ucx)), fabsf(__fsub_rn(hh.y, T.ucy))), T.uext);
                    const float Tf = __fmaf_rn(bc.erel, mu, __fmaf_rn(bc.kabs, __fadd_rn(hh.z, T.ext), bc.abs_floor));
                    const f32x2 hxx = pk(hh.x, hh.x), hyy = pk(hh.y, hh.y);
                    const ulonglong2 *pg = reinterpret_cast<const ulonglong2 *>(T.P) + grp * K3_GSTRIDE + 3 * u0;
                    int sure = 0;
                    uint32_t inband = 0u;
#pragma unroll
                    for (int p = 0; p < PAIRS; ++p) {
                        const ulonglong2 v0 = pg[3 * p], v1 = pg[3 * p + 1], v2 = pg[3 * p + 2];
                        float ep0, ep1, em0, em1;
                        unpk(ffma2(hxx, v0.x, ffma2(hyy, v0.y, v1.x)), ep0, ep1);
                        unpk(ffma2(hxx, v1.y, ffma2(hyy, v2.x, v2.y)), em0, em1);
                        const float m0 = __fsub_rn(ep0, fabsf(em0)), m1 = __fsub_rn(ep1, fabsf(em1));  // a - |tau c|
                        if (!(fabsf(m0) > Tf)) inband |= 1u << (2 * p); else sure += m0 > 0.f;
                        if (!(fabsf(m1) > Tf)) inband |= 2u << (2 * p); else sure += m1 > 0.f;
                    }
                    if (sure) atomicAdd(&R.cnt[hl], sure);
                    // (B) in line: a unit has a few dozen such tests, spread over a few warps -- a second queue, a barrier and a
                    // pass of its own cost more than the divergence here (2 800 -> 2 200 cycles of the producers' 5 500 per unit
                    // at 128 hypotheses, profiles/r02_k_count_tc_timeline.md)
                    while (inband) {
                        const int j = grp * K3_GROUP + 2 * u0 + __ffs(inband) - 1;
                        inband &= inband - 1;
                        if (j >= T.n) continue;  // padded pixels (m = -1e30) only get here when the band is infinite
                        resolve(T, W, hl, j);
                    }
                }
                if (SPLIT) dbar(); else pbar();
                if (gt == 0) TC_STAMP(u + 2, 15);
                if (T.last) {  // the row's last unit: one global add per hypothesis, and the buffer is clean for its next row
                    for (int hl = gt; hl < T.hcount; hl += GN) {
                        const int c = R.cnt[hl];
                        PVV_CHECK(T.cnt_row >= 0 && T.cnt_row + hl < (long long)d.b * d.vn * ws.hn_pad && c <= d.max_num + PVV_CHUNK_PX);
                        if (c) { atomicAdd(&ws.counts[T.cnt_row + hl], c); R.cnt[hl] = 0; }
                    }
                }
            };

            // This CTA's contiguous range of units, in (keypoint, pass, chunk) order; every global load is issued an
            // iteration (or more) before its use: chunk-table entry of unit k+3 | pool offset of unit k+2 | cp.async of
            // the raw inputs of unit k+1 | operands of unit k.
            const int npass = (d.hn + TC_PASS_H - 1) / TC_PASS_H;
            const uint32_t nchunks = (uint32_t)ws.chunk_start[d.b];
            const long long units = (long long)nchunks * d.vn * npass;
            const long long u_lo = units * blockIdx.x / gridDim.x, u_hi = units * (blockIdx.x + 1) / gridDim.x;
            const uint32_t nunits = (uint32_t)(u_hi - u_lo);   // units of this CTA: unit k = global unit u_lo + k
            // (chunk, keypoint-pass group) of unit k, advanced without divisions: unit k + add
            uint32_t ci = u_lo < u_hi ? (uint32_t)(u_lo % nchunks) : 0u, rg = u_lo < u_hi ? (uint32_t)(u_lo / nchunks) : 0u;
            auto unit_at = [&](uint32_t add, uint32_t &co, uint32_t &ro) {
                co = ci + add; ro = rg;
                while (co >= nchunks) { co -= nchunks; ++ro; }
            };
            // the look-ahead loads go through shared memory (cp.async by one thread): a register-carried prefetch would stall
            // on its own load as soon as the loop rotates its registers
            auto fetch_entry = [&](uint32_t add, uint32_t kk) {   // thread 0: entry of unit kk = current unit + add
                if (kk < nunits) {
                    uint32_t co, ro;
                    unit_at(add, co, ro);
                    cp_async16(&S.ent[kk & 7u], ws.chunk_tab + co);
                } else {
                    S.ent[kk & 7u] = make_int4(-1, 0, 0, 0);
                }
            };
            auto fetch_off = [&](uint32_t kk) {                  // thread 0; the entry of unit kk has landed
                const int img = S.ent[kk & 7u].x;
                if (img >= 0) cp_async8(&S.off[kk & 7u], ws.img_off + img);
            };
            // Raw inputs of a unit = three contiguous runs of the pooled SoA buffers (constants, directions, coordinates; a
            // fourth for the hypotheses when the unit opens a row): ONE thread issues them as 1-D bulk copies (TMA:
            // cp.async.bulk, completion counted in bytes on bar_in) instead of 768 cp.async spread over the producers.
            // Sources are 16-byte aligned because every image starts on a 16-pixel boundary of the pool (PVV_POOL_ALIGN); sizes
            // are rounded up to 16 bytes (the pool and a hypothesis row have that slack).
            auto stage_inputs = [&](uint32_t rg1, uint32_t k1, const int4 &e, long long off, bool new_row, int rowbuf) {
                (void)off;
                if (e.x >= 0 && !K3_ABL(1024) && ptid == SPT) {
                    const int vi = (int)rg1 / npass, hbase = ((int)rg1 % npass) * TC_PASS_H;
                    const long long pix = e.y;   // a pool index
                    const int n = e.z & 0x7fff;
                    tc_raw &W = S.raw[k1 % TC_RAWS];
                    unsigned long long *bar = &S.bar_in[k1 % TC_RAWS];
                    const uint32_t b_pq = (uint32_t)n * 16u, b_dir = (uint32_t)((n + 1) & ~1) * 8u, b_xy = (uint32_t)((n + 3) & ~3) * 4u;
                    const int hcount = min(TC_PASS_H, d.hn - hbase);
                    const uint32_t b_h = new_row ? (uint32_t)((hcount + 1) & ~1) * 8u : 0u;
                    tc_expect_tx(bar, b_pq + b_dir + b_xy + b_h);
                    tc_bulk_g2s(W.pq, ws.pix_pq + (long long)vi * ws.cap + pix, b_pq, bar);
                    tc_bulk_g2s(W.dir, ws.pix_dir + (long long)vi * ws.cap + pix, b_dir, bar);
                    tc_bulk_g2s(W.xy, ws.pix_xy + pix, b_xy, bar);
                    if (new_row)
                        tc_bulk_g2s(S.rows[rowbuf].raw, ws.hyp + ((long long)e.x * d.vn + vi) * ws.hn_pad + hbase, b_h, bar);
                }
                cp_async_commit();   // (the look-ahead ring of thread 0 still uses cp.async)
            };
            if (def_role) {
                // ---- deferred-pass group: unit after unit, as soon as the voters are through with it ----
                for (uint32_t u = 0;; ++u) {
                    const uint32_t s = u % TC_STAGES;
                    tc_bar_wait<TC_SLEEP_PROD>(&S.bar_ab[s], (u / TC_STAGES) & 1u);   // the stage's contents are those of unit u
                    if (S.st[s].stop) break;
                    deferred(u);
                    dbar();   // every thread of the group is through with the stage: it may be rebuilt
                    if (lane == 0) tc_bar_arrive(&S.bar_free[s]);
                }
                if (amb_local) atomicAdd((unsigned long long *)&ws.counters[PVV_CTR_AMB_LO], amb_local);
            } else {
            if (ptid == SPT) {
                for (uint32_t i = 0; i < 3; ++i) fetch_entry(i, i);
                cp_async_commit();
                cp_async_wait_all();
                fetch_off(0); fetch_off(1);
                cp_async_commit();
                cp_async_wait_all();
            }
            pbar();
            int4 e0 = S.ent[0];
            int rowbuf = 0, ox = 0, oy = 0;   // row buffer and frame origin of unit k
            bool new_row = true;              // unit k starts a row
            stage_inputs(rg, 0, e0, S.off[0], true, 0);
            uint32_t k = 0;
            for (;; ++k) {
                const uint32_t s = k % TC_STAGES;
                tc_stage &T = S.st[s];  // free: the deferred passes of unit k - 3 ended in the previous iteration
                if (ptid == 0) TC_STAMP(k, 0);
                cp_async_wait_all();
                pbar();  // the look-ahead ring is up to date; everything of the previous iteration is complete
                if (ptid == 0) TC_STAMP(k, 1);
                e0 = S.ent[k & 7u];
                const int4 e1 = S.ent[(k + 1) & 7u];
                // the deferred pass of unit k - 3 is over: stage s, and the raw-input buffer unit k + 1 is staged into, are free
                if (SPLIT && k >= TC_STAGES) tc_bar_wait<TC_SLEEP_PROD>(&S.bar_free[s], (k / TC_STAGES - 1u) & 1u);
                if (e0.x < 0) {
                    if (ptid == 0) T.stop = 1;
                    __syncwarp();
                    if (lane == 0) tc_bar_arrive(&S.bar_ab[s]);
                    break;
                }
                if (!K3_ABL(1024)) tc_bar_wait<TC_SLEEP_VOTER>(&S.bar_in[k % TC_RAWS], (k / TC_RAWS) & 1u);   // unit k's raw inputs landed
                if (ptid == 0) TC_STAMP(k, 2);
                if (ptid == SPT) { fetch_entry(3, k + 3); fetch_off(k + 2); }   // committed with the staging group below
                const long long off1 = S.off[(k + 1) & 7u];
                uint32_t ci1, rg1;
                unit_at(1, ci1, rg1);
                const bool next_new_row = e1.x >= 0 && !(e1.x == e0.x && rg1 == rg);
                const int next_rowbuf = next_new_row ? (rowbuf + 1) % TC_ROWS : rowbuf;
                stage_inputs(rg1, k + 1, e1, off1, next_new_row, next_rowbuf);
                if (ptid == SPT) TC_STAMP(k, 3);
                const tc_raw &W = S.raw[k % TC_RAWS];
                tc_row &R = S.rows[rowbuf];
                const int vi = (int)rg / npass, hbase = ((int)rg % npass) * TC_PASS_H;
                const int hcount = min(TC_PASS_H, d.hn - hbase);
                const int n = e0.z & 0x7fff;
                const int nbn = (n + TC_TILE_PX - 1) / TC_TILE_PX;
                const int mbn = (hcount + 127) / 128;
                if (new_row) {
                    // ---- a new row: frame origin = middle pixel of this chunk; hypotheses -> A rows (once per row) ----
                    ox = (int)((uint32_t)e0.w & 0xffffu); oy = (int)((uint32_t)e0.w >> 16);
                    for (int hl = ptid; hl < mbn * 128 && !K3_ABL(64) && ptid < TC_PB; hl += TC_PB) {
                        const float2 hv = hl < hcount ? R.raw[hl] : make_float2(0.f, 0.f);
                        float hx = __fsub_rn(hv.x, (float)ox), hy = __fsub_rn(hv.y, (float)oy);
                        float mm = __fadd_rn(fabsf(hx), fabsf(hy));
                        if (!(mm < 1e15f)) {  // far-away / non-finite hypothesis: every test takes the exact path
                            hx = 0.f; hy = 0.f; mm = __int_as_float(0x7f800000);
                        }
                        R.H[hl] = make_float4(hx, hy, mm, 0.f);
                        uint32_t xh, xl, yh, yl;
                        tc_split(hx, xh, xl); tc_split(hy, yh, yl);
                        R.A[0][hl] = make_uint4(xh, xh, xl, yh);
                        R.A[1][hl] = make_uint4(yh, yl, 0x3f800000u, 0x3f800000u);
                    }
                }
                // ---- pixel constants in the row's frame: FP32 (phase A) and their TF32 split as B rows ----
                int ext = 0, uext = 0;
                const int ucx = (int)((uint32_t)e0.w & 0xffffu) - ox, ucy = (int)((uint32_t)e0.w >> 16) - oy;  // the unit's centre
#pragma unroll 2
                for (int j = ptid; j < nbn * TC_TILE_PX && !K3_ABL(256) && ptid < TC_PB; j += TC_PB) {
                    px_consts c = never;
                    if (j < n && !K3_ABL(4)) {
                        const float4 pq = W.pq[j];
                        const uint32_t xy = W.xy[j];
                        if (pq.x != 0.f || pq.y != 0.f) {
                            const int ix = (int)(xy & 0xffffu) - ox, iy = (int)(xy >> 16) - oy;
                            const float fx = (float)ix, fy = (float)iy;
                            c.px = pq.x; c.py = pq.y; c.qx = pq.z; c.qy = pq.w;
                            c.pz = -__fmaf_rn(fx, c.px, __fmul_rn(fy, c.py));
                            c.qz = -__fmaf_rn(fx, c.qx, __fmul_rn(fy, c.qy));
                            ext = max(ext, abs(ix) + abs(iy));
                            uext = max(uext, abs(ix - ucx) + abs(iy - ucy));
                        }
                    }
                    k3_store_px(T.P, j, c);
                    uint32_t ph, pl, qh, ql, rh, rl;
                    tc_split(c.px, ph, pl); tc_split(c.py, qh, ql); tc_split(c.pz, rh, rl);
                    T.B[0][2 * j] = make_uint4(ph, pl, ph, qh);
                    T.B[1][2 * j] = make_uint4(ql, qh, rh, rl);
                    tc_split(c.qx, ph, pl); tc_split(c.qy, qh, ql); tc_split(c.qz, rh, rl);
                    T.B[0][2 * j + 1] = make_uint4(ph, pl, ph, qh);
                    T.B[1][2 * j + 1] = make_uint4(ql, qh, rh, rl);
                }
                if (ptid == 0) TC_STAMP(k, 4);
                ext = __reduce_max_sync(0xffffffffu, ext);
                uext = __reduce_max_sync(0xffffffffu, uext);
                if (lane == 0 && ext) { atomicMax(&S.ext, ext); atomicMax(&S.uext, uext); }
                if (!K3_ABL(128)) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // operands visible to the tensor pipe
                pbar();
                if (ptid == 0) {
                    T.stop = 0; T.mbn = mbn; T.nbn = nbn; T.n = n; T.hcount = hcount; T.row = rowbuf; T.qn = 0;
                    T.last = (e1.x < 0 || next_new_row) ? 1 : 0;
                    T.cnt_row = ((long long)e0.x * d.vn + vi) * ws.hn_pad + hbase;
                    T.ext = (float)S.ext; T.uext = (float)S.uext; T.ucx = (float)ucx; T.ucy = (float)ucy;
                    S.ext = 0; S.uext = 0;
                }
                __syncwarp();
                if (lane == 0) tc_bar_arrive(&S.bar_ab[s]);
                if (ptid == 0) TC_STAMP(k, 5);
                // ---- deferred passes of the unit the voters finished last (the other group's job when the roles are split) ----
                if (!SPLIT && k >= 2 && !K3_ABL(512)) deferred(k - 2);
                if (ptid == 0) TC_STAMP(k, 7);
                new_row = next_new_row; rowbuf = next_rowbuf; ci = ci1; rg = rg1;
            }
            if (!SPLIT) {
                // k units were published (0 .. k-1): finish the deferred passes of the last two
                if (k >= 2) deferred(k - 2);
                pbar();
                if (k >= 1) deferred(k - 1);
                if (amb_local) atomicAdd((unsigned long long *)&ws.counters[PVV_CTR_AMB_LO], amb_local);
            }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
#ifdef PVV_TC_PROF
    if (blockIdx.x == 0 && tid == 0) {
        const long long t0 = g_tc_prof[0][0];
        printf("unit: P start | pbar | raw in | staged next | B built | ab arrived | done[k-2] seen | deferred end || V wait ab | V got ab | V votes end | V done || M got ab || V cycles in full_wait | in wait::ld\n");
        for (int u = 0; u < TC_PROF_UNITS; ++u) {
            printf("%3d:", u);
            for (int q = 0; q < 13; ++q) printf(" %7lld", g_tc_prof[u][q] ? g_tc_prof[u][q] - t0 : -1LL);
            printf(" | %6lld %6lld | deferred pass over %7lld\n", g_tc_prof[u][13], g_tc_prof[u][14], g_tc_prof[u][15] - t0);
        }
    }
#endif
    if (warp == TC_MMA_WARP)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TC_TMEM_COLS));
}
