// pvv_vote_tc2.cuh -- K3, tensor-pipe kernel, second generation (included by pvv_vote.cu after pvv_vote_tc.cuh, whose
// helpers -- mbarrier / tcgen05 wrappers, the 3xTF32 split, tc_group -- it shares).
//
// Same arithmetic as k_count_tc (see the header of pvv_vote_tc.cuh: the two linear forms a and tau*c of a test come out
// of one tcgen05.mma kind::tf32 per 128 hypotheses x 128 pixels, 3xTF32 split, accumulators in TMEM; the CUDA cores keep
// m = a - |tau c|, the sign count and the guard band).  What changed is the plumbing around the vote, because the
// round-1 profile showed the vote itself to be 2.3 of 5.66 executed warp instructions per warp-test: 20 % were mbarrier
// polls, the producers' per-unit chain (5 CTA-partial barriers, deferred passes of unit k-2 in the same loop) was as
// long as the vote on a unit, and the voters spilled.
//
//   * The deferred passes LEAVE the kernel.  A flagged (hypothesis, 16-pixel group) pair is appended to a queue in global
//     memory (one slice per CTA, position from one shared-memory atomic per warp and unit) and re-decided by
//     k_count_tc_deferred, a plain full-grid kernel that follows in the PDL chain: FP32 filter on the chunk's own frame
//     (phase A), exact reference sequence inside ITS band (phase B), counts corrected with global atomics.  This kernel
//     no longer keeps FP32 constants, raw directions, hypothesis copies or count arrays in shared memory, and nothing
//     waits for a deferred pass.
//   * Builders are independent WARPS, not a CTA-wide team: builder b owns units b, b+NB, ...; it fetches its next unit's
//     raw inputs with two 1-D bulk copies (cp.async.bulk, complete_tx on its own mbarrier -- TMA), builds the B operand
//     of its unit from shared memory and publishes the stage with one mbarrier arrive.  No bar.sync anywhere in the loop.
//   * The frame of a row is the IMAGE's (origin = its median foreground pixel, written by K1d), so every builder knows it
//     without talking to the builder that opened the row; the A operand is built once per row by whoever owns its first
//     unit.
//   * Every wait is mbarrier.try_wait with a suspend-time hint: a waiting warp is parked by the hardware instead of
//     polling (the polls were 12-20 % of the executed instructions).
//   * Voters keep their hypotheses' inlier counts in REGISTERS across the units of a row and flush them with one global
//     atomic per hypothesis and row.
#pragma once

// development ablations (tools/t2_ablate.py, -DPVV_K3_ABLATE builds only): PVV_K3_ABLATE = bits << 16
//   1 no vote   2 no TMEM loads   4 builders publish without building B   8 no MMA (plain arrive)   16 no queue pushes
//   32 issuer does not wait for free accumulators   64 voters neither wait for tiles nor hand them back
//   128 builders fetch no raw inputs   256 builders build no A rows   512 no tcgen05 fences in the tile loops
#define T2A(bit) K3_ABL((bit) << 16)

constexpr int T2_NS = 6;         // stage ring: B operand + header of units k .. k+5
constexpr int T2_NROW = 3;       // row ring: a row buffer is rewritten 3 rows later
constexpr int T2_VOTER_WARPS = 16;
constexpr int T2_MMA_WARP = 16;                // first builder warp: owns the TMEM allocation (there is no dedicated MMA warp)
constexpr int T2_TILE_PX = 64;                 // pixels per MMA tile
constexpr int T2_TILE_COLS = 2 * T2_TILE_PX;   // N of one MMA = TMEM columns of one accumulator buffer: 4 buffers fill the SM's 512
#ifndef T2_WAIT_MODE
#define T2_WAIT_MODE 0    // 0: try_wait with a suspend-time hint, 1: try_wait (system time limit), 2: test_wait (pure poll)
#endif
constexpr uint32_t T2_SUSPEND_NS = 20000;   // try_wait suspend-time hint; the hardware wakes the warp when the phase completes

// setmaxnreg moves registers inside the CTA's pool, which is what the CTA was LAUNCHED with (threads x launch registers),
// not the SM's whole file (measured on the B200: with a split that needed more, exactly the voters that found the 1024
// registers left over per sub-partition started and the rest waited forever).  ptxas caps the launch count at
// floor(65536 / threads / 8) * 8, so for NB = 4: 640 threads x 96 = 61440 = 512 x 104 + 128 x 64.
// NB must be a multiple of 4: setmaxnreg is executed by whole warpgroups.
template <int NB> struct t2_cfg {
    static_assert(NB == 4, "register split worked out for 4 builder warps");
    static constexpr int threads = (T2_VOTER_WARPS + NB) * 32;
    static constexpr int regs_launch = 96, regs_voter = 104, regs_other = 64;
    static_assert(T2_VOTER_WARPS * 32 * regs_voter + NB * 32 * regs_other == threads * regs_launch, "setmaxnreg pool");
};

struct t2_hdr {                   // written by the unit's builder (lane 0), read by the voters and the MMA issuer
    int stop, n, mbn, nbn, hcount, rowbuf, rowseq, pool0, vi, hbase;   // pool0: pool index of the chunk's first pixel
    uint32_t corg;                // x | y << 16 of the chunk's middle pixel
    float ext;                    // largest |offset| (L1) of a pixel of the unit from the frame origin
    float ucx, ucy, uext;         // the unit's centre in the frame, largest |offset| of its pixels from that centre
    long long cnt_row;            // offset of (image, keypoint) row + pass in counts / hyp
};
struct t2_stage {
    uint4 B[2][2 * PVV_CHUNK_PX]; // pixels x {a, tau c}: [K half][row = 2*pixel + form], canonical K-major core matrices
    t2_hdr h;
};
struct t2_row {
    uint4 A[2][512];              // hypotheses: [K half][row]
    float4 H[512];                // {hx, hy, |hx|+|hy| (inf: every test of this hypothesis takes the exact path), -} in the frame
};
struct t2_raw {                   // raw inputs of one unit, landed by two bulk copies
    float4 pq[PVV_CHUNK_PX];      // frame-independent constants of the pre-pass {nx, ny, tau ny, -tau nx}
    uint32_t xy[PVV_CHUNK_PX];
};
template <int NB> struct t2_smem {
    t2_row rows[T2_NROW];
    t2_stage st[T2_NS];
    t2_raw raw[NB][2];
    unsigned long long bar_ready[T2_NS], bar_done[T2_NS], bar_full[4], bar_empty[4], bar_raw[NB][2];
    uint32_t tmem;
    int qn;                       // entries pushed to this CTA's slice of the deferred queue
};

// mbarrier wait that parks the warp: try_wait with a suspend-time hint returns when the phase completes or the hint
// expires.  The watchdog counts this warp's own failed waits (not wall-clock time: a time-sliced context does not
// advance it); 2^22 expired hints are seconds of continuous waiting where a legitimate wait lasts microseconds.
#ifdef PVV_T2_PROF   // development: cycles each warp spends in each kind of wait, printed by CTA 0 at the end
#define T2_PROF_DECL long long t2_prof_w[8] = {0, 0, 0, 0, 0, 0, 0, 0}; const long long t2_prof_t0 = clock64();
#define T2_WAIT(bar, par, id) do { const long long t2_c0 = clock64(); t2_wait(bar, par, id); t2_prof_w[(id) & 7] += clock64() - t2_c0; } while (0)
#define T2_PROF_PRINT(role) do { if (blockIdx.x == 3 && (threadIdx.x & 31) == 0) printf("prof %s warp %2d total %lld  w1(full) %lld w2(ready) %lld w3(empty) %lld w4(rowbuf) %lld w5(stage) %lld w6(raw) %lld w7 %lld\n", role, threadIdx.x >> 5, clock64() - t2_prof_t0, t2_prof_w[1], t2_prof_w[2], t2_prof_w[3], t2_prof_w[4], t2_prof_w[5], t2_prof_w[6], t2_prof_w[7]); } while (0)
#else
#define T2_PROF_DECL
#define T2_WAIT(bar, par, id) t2_wait(bar, par, id)
#define T2_PROF_PRINT(role)
#endif

__device__ __forceinline__ void t2_wait(uint32_t bar_saddr, uint32_t parity, int id = 0)
{
    uint32_t done;
#ifndef PVV_NO_WATCHDOG
    uint32_t fails = 0;
#endif
    for (;;) {
#if T2_WAIT_MODE == 0
        asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2, %3;\n\tselp.b32 %0, 1, 0, P1;\n\t}"
                     : "=r"(done) : "r"(bar_saddr), "r"(parity), "r"(T2_SUSPEND_NS) : "memory");
#elif T2_WAIT_MODE == 1
        asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.b32 %0, 1, 0, P1;\n\t}"
                     : "=r"(done) : "r"(bar_saddr), "r"(parity) : "memory");
#else
        asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.test_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.b32 %0, 1, 0, P1;\n\t}"
                     : "=r"(done) : "r"(bar_saddr), "r"(parity) : "memory");
#endif
        if (done) return;
#ifdef PVV_T2_DEBUG   // development: report which wait starved and carry on (the kernel then ends with wrong counts)
        if (++fails > (1u << 18)) {
            if ((threadIdx.x & 31) == 0 || id >= 100)
                printf("t2 wait timeout: cta %d warp %d lane %d wait-id %d parity %u\n", blockIdx.x, threadIdx.x >> 5, threadIdx.x & 31, id, parity);
            return;
        }
#elif !defined(PVV_NO_WATCHDOG)
        if (++fails > (1u << 22)) __trap();
#endif
    }
}
__device__ __forceinline__ void t2_wait(unsigned long long *b, uint32_t parity, int id = 0) { t2_wait(tc_saddr(b), parity, id); }

// (the bulk-copy helpers are pvv_vote_tc.cuh's)
__device__ __forceinline__ void t2_bulk_g2s(void *smem, const void *gmem, uint32_t bytes, unsigned long long *bar) { tc_bulk_g2s(smem, gmem, bytes, bar); }
__device__ __forceinline__ void t2_expect_tx(unsigned long long *bar, uint32_t bytes) { tc_expect_tx(bar, bytes); }
__device__ __forceinline__ void t2_fence_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// Entry of the deferred queue, self-contained so that the kernel that resolves it has no dependent table look-ups:
//   .x pool index of the group's first pixel   .y index of the hypothesis in hyp / counts
//   .z pixels of the group that exist (0..16) | keypoint << 8   .w x | y << 16 of the chunk's middle pixel (phase-A frame)
__device__ __forceinline__ uint4 t2_entry(int pool_px, long long hyp_idx, int nvalid, int vi, uint32_t org)
{
    return make_uint4((uint32_t)pool_px, (uint32_t)hyp_idx, (uint32_t)nvalid | ((uint32_t)vi << 8), org);
}

template <int NB>
__global__ void __launch_bounds__(t2_cfg<NB>::threads, 1) k_count_tc2(pvv_dims d, pvv_ws ws, band_consts bc)
{
    constexpr int REGS_VOTER = t2_cfg<NB>::regs_voter, REGS_OTHER = t2_cfg<NB>::regs_other;
    extern __shared__ __align__(128) unsigned char t2_smem_raw[];
    t2_smem<NB> &S = *reinterpret_cast<t2_smem<NB> *>(t2_smem_raw);
    static_assert(K3_GROUP == 16 && PVV_CHUNK_PX == 256, "tile / group geometry");

    pvv_pdl_enter();
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    if (tid == 0) {
        for (int i = 0; i < T2_NS; ++i) { tc_bar_init(&S.bar_ready[i], 1); tc_bar_init(&S.bar_done[i], T2_VOTER_WARPS); }
        for (int i = 0; i < 4; ++i) { tc_bar_init(&S.bar_full[i], 1); tc_bar_init(&S.bar_empty[i], 4); }
        for (int i = 0; i < NB; ++i) { tc_bar_init(&S.bar_raw[i][0], 1); tc_bar_init(&S.bar_raw[i][1], 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        S.qn = 0;
    }
    if (warp == T2_MMA_WARP) {  // the MMA warp owns the TMEM allocation: all 512 columns (four 128-column accumulators)
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tc_saddr(&S.tmem)), "r"(TC_TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = S.tmem;

    if (warp < T2_VOTER_WARPS) {
        // ================= voting warps: TMEM -> registers -> counts =================
        // Warp = (TMEM lane quarter q4 = hypotheses 32*q4 .. +31 of an M-block, phase p): the warps of phase p take the tiles
        // t = p (mod 4), whose accumulator is always TMEM buffer p.  A tile = 128 hypotheses x 64 pixels; a warp pulls its
        // 32 lanes of all four 16-pixel groups.  The four warps of a scheduler are the four phases of one lane quarter, so
        // they work on four different tiles: a buffer is handed back by 4 warps (not 16), and between two tiles of a warp
        // lie three tiles of the others -- the MMA / commit / wake-up latency of the next tile is never exposed.
        asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(REGS_VOTER));
        T2_PROF_DECL
        const int q4 = warp & 3, p = warp >> 2;
        // The MMAs of buffer p are issued by team p itself: its leader warp (the one with q4 == p, so the four leaders sit on
        // four different schedulers) waits until its three siblings have pulled the tile, then issues the team's next tile.
        // A single in-order issuer warp took ~250 cycles per tile (try_wait + issue + commit are a serial latency chain), as
        // long as the four teams vote on a round of four tiles: it, not the vote, bounded the kernel.  Here that latency
        // belongs to one team and is hidden by the other three teams of the scheduler.
        const bool leader = q4 == p;
        uint32_t a_full = tc_saddr(&S.bar_full[p]), a_empty = tc_saddr(&S.bar_empty[p]);
        uint32_t tmem_row = tmem + ((uint32_t)(q4 * 32) << 16) + (uint32_t)(p * T2_TILE_COLS);
        uint32_t lane0 = lane == 0;
        asm volatile("" : "+r"(a_full), "+r"(a_empty), "+r"(tmem_row), "+r"(lane0));  // kept in registers, not rematerialised
        // Tile slots of a unit are numbered j = nb * S + mb (nb: 64-pixel quarter of the chunk, mb: 128-hypothesis M-block,
        // S = 1, 2, 4, 4 for 1, 2, 3, 4 M-blocks; with 3 the slots mb = 3 stay empty) and slot t of the running count uses
        // buffer t & 3.  S divides 4, so within a unit -- and, as long as the chunks are full, within a whole row -- a warp
        // sees ONE M-block, i.e. one hypothesis per thread: its band half-width is computed once per unit, its inlier count
        // is one register, the sign lanes are summed once per unit.
        uint32_t tb = 0;           // slot count at the start of the unit, in lock-step with the MMA issuer's
        uint32_t used = 0;         // tiles this warp has taken from its buffer (parity of the next `full` phase)
        int cnt = 0;               // inlier count of this thread's hypothesis, accumulated over the units of a row
        int cur_seq = -1, cur_mb = -1;
        long long cur_idx = -1;    // counts index of that hypothesis (-1: none / beyond hcount)
        unsigned long long amb_local = 0;
        uint4 *slice = ws.tcq + (size_t)blockIdx.x * ws.tcq_slice_cap;

        auto flush = [&] {  // the row (or the warp's M-block) is over: one global add per hypothesis
            if (cnt && cur_idx >= 0) atomicAdd(&ws.counts[cur_idx], cnt);
            cnt = 0;
        };
        auto landed = [&] { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); };

        // issue cursor of the leader: walks the same (unit, slot) sequence as the vote loop, one tile ahead
        // kind::tf32: D = f32 (bit 4), A = B = tf32 (2 << 7, 2 << 10), K-major both, N >> 3 at bit 17, M >> 4 at bit 24
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(T2_TILE_COLS >> 3) << 17) | ((128u >> 4) << 24);
        uint32_t is = 0, iph = 0, itb = 0, n_iss = 0, iunit = 0, vunit = 0;
        int ij = -1;           // next slot of this team in the unit the cursor is in (-1: not positioned yet)
        bool idone = false;
        // Issues the team's next tile if none is in flight.  Called before the loop, at the top of every unit and after every
        // release.  The cursor never runs more than 2 units ahead of the vote loop: the builder of unit k waits, when k opens a
        // row, until the voters are done with unit k - 3 (row ring), so units beyond vunit + 2 may depend on this very warp's
        // progress.  A team without tiles (3 M-blocks) just trails along.
        auto issue_next = [&] {
            while (!idone && n_iss == used && iunit <= vunit + 2) {
                t2_stage &U = S.st[is];
                T2_WAIT(&S.bar_ready[is], iph, 4);
                if (U.h.stop) { idone = true; break; }
                const int umbn = U.h.mbn, ush = umbn >= 3 ? 2 : umbn - 1, unslots = U.h.nbn << ush;
                const int um0 = (int)(((uint32_t)p - itb) & 3u), umb = um0 & ((1 << ush) - 1);
                if (ij < 0) ij = um0;
                if (umb < umbn && ij < unslots) {
                    // the n-th tile of the buffer (n >= 1) needs the release of its (n-1)-th by all four warps of the team
                    if (n_iss && !T2A(64)) T2_WAIT(a_empty, (n_iss - 1u) & 1u, 3);
                    ++n_iss;
                    if (!T2A(512)) tc_fence_after();
                    // descriptors advance linearly: 128 rows x 16 B = 2048 B = 128 units of 16 B in the start-address field
                    const uint64_t da = tc_desc(tc_saddr(&S.rows[U.h.rowbuf].A[0][0]), sizeof(S.rows[0].A[0]), 128) + (uint64_t)(umb * 128);
                    const uint64_t db = tc_desc(tc_saddr(&U.B[0][0]), sizeof(U.B[0]), 128) + (uint64_t)((ij >> ush) * 128);
                    if (T2A(8)) {
                        if (lane == 0) tc_bar_arrive(a_full);
                    } else if (tc_elect_one()) {
                        asm volatile(
                            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                            "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem + (uint32_t)(p * T2_TILE_COLS)),
                            "l"(da), "l"(db), "r"(idesc), "r"(0));
                        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(a_full) : "memory");
                    }
                    __syncwarp();
                    ij += 4;
                    break;
                }
                itb += (uint32_t)unslots;   // no (further) slot of this team in the unit: on to the next one
                ij = -1;
                ++iunit;
                if (++is == T2_NS) { is = 0; iph ^= 1u; }
            }
        };
        if (leader) issue_next();

        uint32_t s = 0, ph = 0;
        for (;; ++vunit) {
            t2_stage &T = S.st[s];
            T2_WAIT(&S.bar_ready[s], ph, 2);
            if (T.h.stop) break;
            if (leader) issue_next();
            const int mbn = T.h.mbn, nbn = T.h.nbn, hcount = T.h.hcount;
            const int sh = mbn >= 3 ? 2 : mbn - 1;           // log2(S)
            const int m0 = (int)(((uint32_t)p - tb) & 3u);   // this warp's first slot of the unit
            const int mb = m0 & ((1 << sh) - 1);
            const int nslots = nbn << sh;
            tb += (uint32_t)nslots;
            if (mb < mbn) {
                const int hl = mb * 128 + q4 * 32 + lane;
                if (T.h.rowseq != cur_seq || mb != cur_mb) {
                    flush();
                    cur_seq = T.h.rowseq; cur_mb = mb;
                    cur_idx = hl < hcount ? T.h.cnt_row + hl : -1;
                }
                const t2_row &R = S.rows[T.h.rowbuf];
                const float4 hh = R.H[hl];
                const float mu = __fadd_rn(__fadd_rn(fabsf(__fsub_rn(hh.x, T.h.ucx)), fabsf(__fsub_rn(hh.y, T.h.ucy))), T.h.uext);
                const float Tt = __fmaf_rn(bc.erel, mu, __fmaf_rn(bc.kabs_tc, __fadd_rn(hh.z, T.h.ext), bc.abs_floor));
                int neg = 0, ktile = 0;
                uint32_t dmask = 0u;   // bit 4k + g: flagged 16-pixel group g of the k-th tile this warp took from the unit
                uint32_t v0[32], v1[32];
#pragma unroll 1
                for (int j = m0; j < nslots; j += 4, ++ktile) {
                    uint32_t dm = 0u;
                    if (!T2A(64)) T2_WAIT(a_full, used & 1u, 1);
                    ++used;
                    if (!T2A(512)) tc_fence_after();
                    const bool ld = !T2A(2), vt = !T2A(1);
                    if (ld) { tc_ld32(tmem_row, v0); tc_ld32(tmem_row + 32u, v1); landed(); }
                    if (vt) tc_group(v0, Tt, neg, dm, 1u);
                    if (ld) tc_ld32(tmem_row + 64u, v0);
                    if (vt) tc_group(v1, Tt, neg, dm, 2u);
                    if (ld) { tc_ld32(tmem_row + 96u, v1); landed(); }
                    if (!T2A(512)) tc_fence_before();          // the whole tile is in registers: its TMEM buffer is free
                    if (lane0 && !T2A(64)) tc_bar_arrive(a_empty);
                    if (vt) tc_group(v0, Tt, neg, dm, 4u);
                    if (leader) issue_next();   // the siblings have released the buffer by now (same tile, other schedulers)
                    if (vt) tc_group(v1, Tt, neg, dm, 8u);
                    else neg += v0[0] ^ v1[31];
                    dmask |= dm << (4 * ktile);
                }
                if (T2A(16) || cur_idx < 0) dmask = 0u;
                if (cur_idx >= 0) cnt += K3_GROUP * (4 * ktile - __popc(dmask)) - sign_lanes_total(neg);
                // queue the flagged (hypothesis, group) pairs for k_count_tc_deferred: one reservation per warp and unit
                const int mine = __popc(dmask);
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
                    const int pool0 = T.h.pool0, vi = T.h.vi, n = T.h.n;
                    const uint32_t corg = T.h.corg;
                    const int nb0 = m0 >> sh, nbstep = 4 >> sh;
                    while (dmask) {
                        const int bitn = __ffs(dmask) - 1;
                        dmask &= dmask - 1;
                        const int grp = (nb0 + (bitn >> 2) * nbstep) * 4 + (bitn & 3);  // 16-pixel group within the chunk
                        if (q < ws.tcq_slice_cap) {
                            slice[q] = t2_entry(pool0 + grp * K3_GROUP, cur_idx, max(0, min(K3_GROUP, n - grp * K3_GROUP)), vi, corg);
                        } else {  // slice full (pathological inputs): the reference's own decision for the whole group, in line
                            const float2 hv = ws.hyp[cur_idx];
                            int c = 0;
                            for (int jj = grp * K3_GROUP; jj < min(n, (grp + 1) * K3_GROUP); ++jj) {
                                const float2 dir = ws.pix_dir[(long long)vi * ws.cap + pool0 + jj];
                                const uint32_t xy = ws.pix_xy[pool0 + jj];
                                c += pvv_exact_test(dir.x, dir.y, pvv_norm1(dir.x, dir.y), (float)(xy & 0xffffu), (float)(xy >> 16),
                                                    hv.x, hv.y, d.thresh) ? 1 : 0;
                                ++amb_local;
                            }
                            cnt += c;
                        }
                        ++q;
                    }
                }
            }
            __syncwarp();
            if (lane == 0) tc_bar_arrive(&S.bar_done[s]);
            if (++s == T2_NS) { s = 0; ph ^= 1u; }
        }
        flush();
        T2_PROF_PRINT("voter  ");
        if (amb_local) atomicAdd((unsigned long long *)&ws.counters[PVV_CTR_AMB_LO], amb_local);
    } else {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(REGS_OTHER));
        T2_PROF_DECL
        {
            // ================= builders: warp b owns units b, b + NB, ... of this CTA's range =================
            const int bw = warp - T2_VOTER_WARPS;
            // This CTA's contiguous range of units, in (keypoint, pass, chunk) order: unit k = global unit u_lo + k
            const uint32_t npass = (uint32_t)(d.hn + TC_PASS_H - 1) / TC_PASS_H;
            const uint32_t nchunks = (uint32_t)ws.chunk_start[d.b];
            const uint32_t nimg = ws.counters[PVV_CTR_NIMG];
            const unsigned long long units = (unsigned long long)nchunks * d.vn * npass;
            const unsigned long long u_lo = units * blockIdx.x / gridDim.x, u_hi = units * (blockIdx.x + 1) / gridDim.x;
            const uint32_t nunits = (uint32_t)(u_hi - u_lo);
            // (chunk, keypoint-pass group) of unit k
            auto locate = [&](uint32_t k, uint32_t &ci, uint32_t &rg) {
                const unsigned long long u = u_lo + k;
                rg = (uint32_t)(u / nchunks);
                ci = (uint32_t)(u - (unsigned long long)rg * nchunks);
            };
            const int4 none = make_int4(-1, 0, 0, 0);
            auto entry_of = [&](uint32_t k) -> int4 {
                if (k >= nunits) return none;
                uint32_t ci, rg;
                locate(k, ci, rg);
                return __ldg(&ws.chunk_tab[ci]);
            };
            // raw inputs of unit k (entry e) -> raw slot `slot` of this builder: two bulk copies, lane 0
            auto fetch_raw = [&](uint32_t k, const int4 &e, int slot) {
                if (e.x < 0 || lane != 0 || T2A(128)) return;
                uint32_t ci, rg;
                locate(k, ci, rg);
                const int vi = (int)(rg / npass), n = e.z & 0x7fff;
                const uint32_t bytes_pq = (uint32_t)n * 16u, bytes_xy = (uint32_t)((n + 3) & ~3) * 4u;
                t2_raw &W = S.raw[bw][slot];
                t2_expect_tx(&S.bar_raw[bw][slot], bytes_pq + bytes_xy);
                t2_bulk_g2s(W.pq, ws.pix_pq + (long long)vi * ws.cap + e.y, bytes_pq, &S.bar_raw[bw][slot]);
                t2_bulk_g2s(W.xy, ws.pix_xy + e.y, bytes_xy, &S.bar_raw[bw][slot]);
            };
            // row sequence number of the CTA's first unit (rows are numbered (keypoint-pass group) * images + image rank)
            uint32_t seq0 = 0;
            if (nunits) {
                uint32_t ci, rg;
                locate(0, ci, rg);
                seq0 = rg * nimg + ((uint32_t)__ldg(&ws.chunk_tab[ci]).z >> 16);
            }

            int4 eA = entry_of((uint32_t)bw), eB = entry_of((uint32_t)bw + NB);
            uint32_t orgA = eA.x >= 0 ? __ldg(&ws.img_org[eA.x]) : 0u;
            fetch_raw((uint32_t)bw, eA, 0);
            uint32_t it = 0;
            for (uint32_t k = (uint32_t)bw; k < nunits; k += NB, ++it) {
                const uint32_t s = k % T2_NS, kd = k / T2_NS;
                const int slot = (int)(it & 1u);
                // look-ahead: entry of unit k + 2 NB, frame origin and raw inputs of unit k + NB
                const int4 eC = entry_of(k + 2 * NB);
                const uint32_t orgB = eB.x >= 0 ? __ldg(&ws.img_org[eB.x]) : 0u;
                fetch_raw(k + NB, eB, slot ^ 1);

                uint32_t ci, rg;
                locate(k, ci, rg);
                const int vi = (int)(rg / npass), hbase = (int)(rg % npass) * TC_PASS_H;
                const int hcount = min(TC_PASS_H, d.hn - hbase);
                const int n = eA.z & 0x7fff;
                const int nbn = (n + T2_TILE_PX - 1) / T2_TILE_PX, mbn = (hcount + 127) / 128;
                const uint32_t rowseq = rg * nimg + ((uint32_t)eA.z >> 16) - seq0;
                const int rowbuf = (int)(rowseq % T2_NROW);
                const bool new_row = k == 0 || ((eA.z >> 15) & 1);
                const int ox = (int)(orgA & 0xffffu), oy = (int)(orgA >> 16);
                const long long hrow = ((long long)eA.x * d.vn + vi) * ws.hn_pad + hbase;

                if (new_row) {
                    // the row buffer's previous occupant (3 rows back) ended with a unit <= k - 3: the voters (in order) are past it
                    if (k >= 3) T2_WAIT(&S.bar_done[(k - 3) % T2_NS], ((k - 3) / T2_NS) & 1u, 4);
                    t2_row &R = S.rows[rowbuf];
                    for (int hl = lane; hl < mbn * 128 && !T2A(256); hl += 32) {
                        const float2 hv = hl < hcount ? ws.hyp[hrow + hl] : make_float2(0.f, 0.f);
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
                // stage s is free once the voters are done with unit k - NS
                if (k >= T2_NS) T2_WAIT(&S.bar_done[s], (kd - 1u) & 1u, 5);
                if (!T2A(128)) T2_WAIT(&S.bar_raw[bw][slot], (it >> 1) & 1u, 6);   // raw inputs of unit k have landed
                t2_stage &T = S.st[s];
                const t2_raw &W = S.raw[bw][slot];
                // ---- pixel constants in the image's frame, TF32-split, as B rows ----
                int ext = 0, uext = 0;
                const int ucx = (int)((uint32_t)eA.w & 0xffffu) - ox, ucy = (int)((uint32_t)eA.w >> 16) - oy;  // the unit's centre
                for (int j = lane; j < nbn * T2_TILE_PX && !T2A(4); j += 32) {
                    float px = 0.f, py = 0.f, pz = -1e30f, qx = 0.f, qy = 0.f, qz = -1e30f;  // m = -2e30: never an inlier, never in the band
                    if (j < n) {
                        const float4 pq = W.pq[j];
                        const uint32_t xy = W.xy[j];
                        if (pq.x != 0.f || pq.y != 0.f) {
                            const int ix = (int)(xy & 0xffffu) - ox, iy = (int)(xy >> 16) - oy;
                            const float fx = (float)ix, fy = (float)iy;
                            px = pq.x; py = pq.y; qx = pq.z; qy = pq.w;
                            pz = -__fmaf_rn(fx, px, __fmul_rn(fy, py));
                            qz = -__fmaf_rn(fx, qx, __fmul_rn(fy, qy));
                            ext = max(ext, abs(ix) + abs(iy));
                            uext = max(uext, abs(ix - ucx) + abs(iy - ucy));
                        }
                    }
                    uint32_t ph_, pl, qh, ql, rh, rl;
                    tc_split(px, ph_, pl); tc_split(py, qh, ql); tc_split(pz, rh, rl);
                    T.B[0][2 * j] = make_uint4(ph_, pl, ph_, qh);
                    T.B[1][2 * j] = make_uint4(ql, qh, rh, rl);
                    tc_split(qx, ph_, pl); tc_split(qy, qh, ql); tc_split(qz, rh, rl);
                    T.B[0][2 * j + 1] = make_uint4(ph_, pl, ph_, qh);
                    T.B[1][2 * j + 1] = make_uint4(ql, qh, rh, rl);
                }
                ext = __reduce_max_sync(0xffffffffu, ext);
                uext = __reduce_max_sync(0xffffffffu, uext);
                if (lane == 0) {
                    t2_hdr &h = T.h;
                    h.stop = 0; h.n = n; h.mbn = mbn; h.nbn = nbn; h.hcount = hcount; h.rowbuf = rowbuf; h.rowseq = (int)rowseq;
                    h.pool0 = eA.y; h.vi = vi; h.hbase = hbase; h.corg = (uint32_t)eA.w;
                    h.ext = (float)ext; h.ucx = (float)ucx; h.ucy = (float)ucy; h.uext = (float)uext;
                    h.cnt_row = hrow;
                }
                t2_fence_async();   // operands (generic-proxy stores) visible to the tensor pipe; raw slot reads precede its next bulk copy
                __syncwarp();
                if (lane == 0) tc_bar_arrive(&S.bar_ready[s]);
                eA = eB; eB = eC; orgA = orgB;
            }
            // the builder that would own unit `nunits` publishes the stop marker in its stage
            if (nunits % NB == (uint32_t)bw) {
                const uint32_t k = nunits, s = k % T2_NS;
                if (k >= T2_NS) T2_WAIT(&S.bar_done[s], (k / T2_NS - 1u) & 1u, 7);
                if (lane == 0) { S.st[s].h.stop = 1; tc_bar_arrive(&S.bar_ready[s]); }
            }
            T2_PROF_PRINT("builder");
            __syncwarp();
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == T2_MMA_WARP)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TC_TMEM_COLS));
    if (tid == 0) ws.tcq_cnt[blockIdx.x] = S.qn;
}

// ---- K3c: the deferred passes of the tensor-pipe kernel, as a kernel of their own ---------------------------------
// One queue entry = one flagged (hypothesis, 16-pixel group): the tensor-pipe value of at least one of its tests was
// inside the tensor-pipe band, so k_count_tc2 counted NONE of its 16 tests.  SIXTEEN lanes per entry, one pixel each, so a
// warp's loads are two contiguous 256-byte runs (the first version gave every lane 8 pixels: 32 different cache lines per
// load instruction, and the kernel was bound by the load unit at 170 us):
//   phase A  the FP32 filter in the chunk's own frame (origin = the chunk's middle pixel, the frame of the FP32-pipe
//            kernel): a = fma(hx, nx, fma(hy, ny, ra)), tau c = fma(hx, tau ny, fma(hy, -tau nx, rc)), m = a - |tau c| on
//            the pre-pass constants, r formed in fp32 (one rounded product, one fma).  |m| > T decides (bound: DESIGN.md 4a,
//            T = krel * (|hx| + |hy| + extent of the group's pixels) + floor, the FP32-pipe kernel's band);
//   phase B  the tests inside that band get the reference's exact rounding sequence (pvv_exact_test), densely packed.
// Grid: blockIdx.x = slice + tcq_slices * part; `parts` CTAs stride over the entries of one slice.
constexpr int T2_DEF_THREADS = 256;
__global__ void __launch_bounds__(T2_DEF_THREADS) k_count_tc_deferred(pvv_dims d, pvv_ws ws, band_consts bc, int parts)
{
    constexpr int EPB = T2_DEF_THREADS / K3_GROUP;         // entries per CTA and step
    __shared__ uint32_t s_q2[T2_DEF_THREADS];              // phase B: (local entry << 8) | pixel within the group
    __shared__ int s_q2n;
    pvv_pdl_enter();
    const int tid = threadIdx.x, px = tid & (K3_GROUP - 1), le = tid / K3_GROUP;
    const int slice = blockIdx.x % ws.tcq_slices, part = blockIdx.x / ws.tcq_slices;
    const int total = min(ws.tcq_cnt[slice], ws.tcq_slice_cap);
    const uint4 *q = ws.tcq + (size_t)slice * ws.tcq_slice_cap;
    const uint32_t half_mask = 0xffffu << (16 * ((tid >> 4) & 1));   // the 16 lanes of this entry within the warp
    unsigned long long amb_local = 0;
    for (int base = part * EPB; base < total; base += parts * EPB) {
        if (tid == 0) s_q2n = 0;
        __syncthreads();
        const int ei = base + le;
        if (ei < total) {   // uniform over the 16 lanes of an entry
            const uint4 en = q[ei];
            const int nvalid = (int)(en.z & 0xffu), vi = (int)(en.z >> 8);
            const bool ok = px < nvalid;
            const float4 c = ok ? ws.pix_pq[(long long)vi * ws.cap + en.x + px] : make_float4(0.f, 0.f, 0.f, 0.f);
            const uint32_t xy = ok ? ws.pix_xy[en.x + px] : 0u;
            const float2 hv = ws.hyp[en.y];
            const int ox = (int)(en.w & 0xffffu), oy = (int)(en.w >> 16);
            float hx = __fsub_rn(hv.x, (float)ox), hy = __fsub_rn(hv.y, (float)oy);
            const float mm = __fadd_rn(fabsf(hx), fabsf(hy));
            const bool far = !(mm < 1e15f);
            if (far) { hx = 0.f; hy = 0.f; }
            float m = -1e30f;   // padded pixel / pixel the reference never counts: not an inlier, not in the band
            int ext = 0;
            if (c.x != 0.f || c.y != 0.f) {
                const int ix = (int)(xy & 0xffffu) - ox, iy = (int)(xy >> 16) - oy;
                const float fx = (float)ix, fy = (float)iy;
                const float ra = -__fmaf_rn(fx, c.x, __fmul_rn(fy, c.y));
                const float rc = -__fmaf_rn(fx, c.z, __fmul_rn(fy, c.w));
                const float a = __fmaf_rn(hx, c.x, __fmaf_rn(hy, c.y, ra));
                const float tc = __fmaf_rn(hx, c.z, __fmaf_rn(hy, c.w, rc));
                m = __fsub_rn(a, fabsf(tc));
                ext = abs(ix) + abs(iy);
            }
            // the group's extent: maximum over its 16 lanes
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) ext = max(ext, __shfl_xor_sync(half_mask, ext, o, 16));
            const float T = far ? __int_as_float(0x7f800000) : __fmaf_rn(bc.krel, __fadd_rn(mm, (float)ext), bc.abs_floor);
            const bool decided = fabsf(m) > T;
            const uint32_t sure = __ballot_sync(half_mask, decided && m > 0.f) & half_mask;
            if (px == 0 && sure) atomicAdd(&ws.counts[en.y], __popc(sure));
            if (!decided && ok)     // (padded pixels only get here when the band is infinite)
                s_q2[atomicAdd(&s_q2n, 1)] = ((uint32_t)le << 8) | (uint32_t)px;
        }
        __syncthreads();
        const int q2 = s_q2n;
        for (int idx = tid; idx < q2; idx += T2_DEF_THREADS) {
            const uint32_t ent = s_q2[idx];
            const uint4 en = q[base + (int)(ent >> 8)];
            const int j = (int)(ent & 0xffu), vi = (int)(en.z >> 8);
            const float2 hv = ws.hyp[en.y];
            const float2 dir = ws.pix_dir[(long long)vi * ws.cap + en.x + j];
            const uint32_t xy = ws.pix_xy[en.x + j];
            if (pvv_exact_test(dir.x, dir.y, pvv_norm1(dir.x, dir.y), (float)(xy & 0xffffu), (float)(xy >> 16), hv.x, hv.y, d.thresh))
                atomicAdd(&ws.counts[en.y], 1);
            ++amb_local;
        }
        __syncthreads();
    }
    if (amb_local) atomicAdd((unsigned long long *)&ws.counters[PVV_CTR_AMB_LO], amb_local);
}
