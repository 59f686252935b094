// pvv_compact.cu -- K1: order-preserving stream compaction of foreground pixels (HBM-bound).
//
// Replaces, for the whole batch at once and without a host round-trip,
//   cur_mask = mask[bi].byte(); foreground_num = sum(cur_mask)            ransac_voting_gpu.py:533-534
//   (or cur_mask = mask[bi] == 1                                          ransac_voting_gpu.py:339-340)
//   skip if foreground < min_num                                          :537 / :343
//   Bernoulli(max_num/foreground) subsample when foreground > max_num     :543-546 / :351-355
//   coords = nonzero(cur_mask)[:, [1,0]];  direct = vertex[bi].masked_select(...)   :548-551 / :357-360
// The output is a pooled SoA buffer (pix_xy[cap], pix_dir[vn][cap]) in the same row-major order
// torch.nonzero produces, so pixel index t of image i here == row t of the reference's coords/direct.
//
// Four launches: tile counts + foreground bitmask (the only pass over the mask) -> (optional) Bernoulli selection
// applied to the bitmask of images above max_num -> per-image scan + batch scan by the last block -> scatter from the
// bitmask.  The mask is read ONCE through its real strides and dtype; the vertex tensor is read through its real
// strides (the network emits a permuted view), only at foreground pixels.
#include "pvv_internal.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kPxPerThread = PVV_TILE_PX / kThreads;  // 4

__device__ __forceinline__ int warp_incl_scan(int v)
{
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) >= o) v += n;
    }
    return v;
}

// exclusive scan over the block's kThreads values; returns this thread's prefix, *total = block sum
__device__ __forceinline__ int block_excl_scan(int v, int *total)
{
    __shared__ int s_warp[kThreads / 32];
    __shared__ int s_total;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int inc = warp_incl_scan(v);
    if (lane == 31) s_warp[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        int w = lane < kThreads / 32 ? s_warp[lane] : 0;
        const int winc = warp_incl_scan(w);
        if (lane < kThreads / 32) s_warp[lane] = winc - w;
        if (lane == kThreads / 32 - 1) s_total = winc;
    }
    __syncthreads();
    const int r = s_warp[wid] + inc - v;
    *total = s_total;
    __syncthreads();  // s_warp / s_total are reused by the next call
    return r;
}

__device__ __forceinline__ float sel_value(const pvv_dims &d, const float *selection, int img, uint32_t p)
{
    if (selection) return selection[(long long)img * d.h * d.w + p];
    return pvv_draw_uniform(d.seed, d.image_id_base + img, p);
}

// ---- K1a: the ONE pass over the mask: foreground bitmask + per-tile count + per-image foreground sum ----------
// Thread t of the CTA reads pixels t, t+256, t+512, t+768 of the tile (fully coalesced for a contiguous mask of any
// dtype); one ballot per warp and step yields a 32-pixel word of the bitmask in pixel order.
// DT: mask element type (compile-time: no per-element dispatch); LINEAR: rows are contiguous and adjacent, so the
// element offset is img*stride0 + pixel (no division).
template <int DT>
__device__ __forceinline__ uint32_t mask_value_t(const pvv_dims &d, const void *mask, long long off, int fg_rule)
{
    long long v;
    if (DT == PVV_MASK_LOGITS_F32) {
        // torch.argmax(seg_pred, 1) fused into the pass: first maximal class, a NaN counts as maximal (ATen's rule)
        const float *p = (const float *)mask + off;
        float best = p[0];
        int bi = 0;
        for (int c = 1; c < d.seg_c; ++c) {
            const float x = p[c * d.seg_cs];
            if (x > best || (x != x && best == best)) { best = x; bi = c; }
        }
        v = bi;
    }
    else if (DT == PVV_MASK_U8) v = ((const uint8_t *)mask)[off];
    else if (DT == PVV_MASK_I32) v = ((const int32_t *)mask)[off];
    else if (DT == PVV_MASK_I64) v = ((const long long *)mask)[off];
    else {  // float: .byte() truncates toward zero, == 1 compares the float
        const float f = ((const float *)mask)[off];
        if (fg_rule == PVV_FG_EQ1) return f == 1.0f ? 1u : 0u;
        v = (long long)f;
    }
    if (fg_rule == PVV_FG_EQ1) return v == 1 ? 1u : 0u;
    return (uint32_t)(v & 0xFF);
}

template <int DT, bool LINEAR>
__global__ void __launch_bounds__(kThreads) k_tile_count(pvv_dims d, const void *mask, pvv_ws ws)
{
    pvv_pdl_enter();  // first kernel of the chain (plain launch after the memset): the wait is a no-op, the trigger is not
    const int img = blockIdx.y, tile = blockIdx.x;
    const int npx = d.h * d.w;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t *bits = ws.fg_bits + ((long long)img * ws.ntiles + tile) * (PVV_TILE_PX / 32);
    uint32_t v[kPxPerThread];
#pragma unroll
    for (int k = 0; k < kPxPerThread; ++k) {  // all loads first: 4 independent requests in flight per thread
        const int p = tile * PVV_TILE_PX + k * kThreads + threadIdx.x;
        v[k] = 0;
        if (p < npx) {
            long long off;
            if (LINEAR) off = img * d.ms[0] + p;
            else { const int y = p / d.w, x = p - y * d.w; off = img * d.ms[0] + y * d.ms[1] + x * d.ms[2]; }
            v[k] = mask_value_t<DT>(d, mask, off, d.fg_rule);
        }
    }
    int n = 0, s = 0;
#pragma unroll
    for (int k = 0; k < kPxPerThread; ++k) {
        const uint32_t word = __ballot_sync(0xffffffffu, v[k] != 0u);
        if (lane == 0) { bits[k * (kThreads / 32) + wid] = word; n += __popc(word); }
        s += (int)v[k];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    __shared__ int s_n[kThreads / 32], s_s[kThreads / 32];
    if (lane == 0) { s_n[wid] = n; s_s[wid] = s; }
    __syncthreads();
    if (threadIdx.x == 0) {
        int tn = 0, ts = 0;
#pragma unroll
        for (int i = 0; i < kThreads / 32; ++i) { tn += s_n[i]; ts += s_s[i]; }
        ws.tile_cnt[img * ws.ntiles + tile] = tn;
        if (ts) atomicAdd((unsigned long long *)&ws.img_fg[img], (unsigned long long)ts);
    }
}

template <int DT>
static void launch_tile_count(const pvv_dims &d, const void *mask, const pvv_ws &ws, dim3 grid, cudaStream_t st)
{
    const bool linear = d.ms[2] == 1 && d.ms[1] == d.w;
    pvv_hint_carveout(linear ? (const void *)k_tile_count<DT, true> : (const void *)k_tile_count<DT, false>);
    if (linear) k_tile_count<DT, true><<<grid, kThreads, 0, st>>>(d, mask, ws);
    else k_tile_count<DT, false><<<grid, kThreads, 0, st>>>(d, mask, ws);
}

// ---- K1a': images above max_num: apply the Bernoulli selection to the bitmask, recount their tiles ------------
// One warp per tile (lane = one 32-pixel word); a CTA whose image is not subsampled exits at once.
__global__ void __launch_bounds__(kThreads) k_tile_recount(pvv_dims d, const float *selection, pvv_ws ws)
{
    pvv_pdl_enter();
    const int img = blockIdx.y;
    const long long fg = ws.img_fg[img];
    if (fg <= d.max_num || fg < d.min_num) return;
    const int tile = blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (tile >= ws.ntiles) return;
    const float prob = pvv_subsample_prob(d.max_num, fg);
    uint32_t *bits = ws.fg_bits + ((long long)img * ws.ntiles + tile) * (PVV_TILE_PX / 32);
    uint32_t word = bits[lane], keep = 0u;
    while (word) {
        const int bit = __ffs(word) - 1;
        word &= word - 1;
        if (sel_value(d, selection, img, (uint32_t)(tile * PVV_TILE_PX + lane * 32 + bit)) < prob) keep |= 1u << bit;
    }
    bits[lane] = keep;
    int n = __popc(keep);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
    if (lane == 0) ws.tile_cnt[img * ws.ntiles + tile] = n;
}

// ---- K1c: per-image exclusive scan of tile counts; the last block scans the batch ------------------
__global__ void __launch_bounds__(kThreads) k_scan(pvv_dims d, pvv_ws ws)
{
    pvv_pdl_enter();
    const int img = blockIdx.x;
    const long long fg = ws.img_fg[img];
    int flags = 0;
    if (fg < d.min_num) flags |= PVV_IMG_SKIPPED;
    else if (fg > d.max_num) flags |= PVV_IMG_SUBSAMPLED;

    int carry = 0;
    for (int base = 0; base < ws.ntiles; base += kThreads) {
        const int t = base + threadIdx.x;
        const int v = (t < ws.ntiles && !(flags & PVV_IMG_SKIPPED)) ? ws.tile_cnt[img * ws.ntiles + t] : 0;
        int total;
        const int ex = block_excl_scan(v, &total);
        if (t < ws.ntiles) ws.tile_off[img * ws.ntiles + t] = carry + ex;
        carry += total;
    }
    __shared__ bool s_last;
    if (threadIdx.x == 0) {
        ws.tn[img] = carry;
        ws.img_flags[img] = flags;
        if (carry) atomicAdd((unsigned long long *)&ws.counters[PVV_CTR_NEED_LO], (unsigned long long)pvv_pool_span(carry));
        __threadfence();
        s_last = atomicAdd(&ws.counters[PVV_CTR_TICKET], 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();

    // Pool overflow (a user-set pool_capacity that is too small, or an injected `selection` that keeps far more than
    // max_num pixels): images are admitted greedily in batch order; one that does not fit is dropped (zero keypoints,
    // PVV_IMG_OVERFLOW) and takes no room, so every later image that still fits is kept.  Sequential, but b is small and
    // this only runs when the batch as a whole does not fit (every image's block added its span to PVV_CTR_NEED).
    if (*(volatile unsigned long long *)&ws.counters[PVV_CTR_NEED_LO] > (unsigned long long)ws.cap) {
        if (threadIdx.x == 0) {
            long long used = 0;
            for (int i = 0; i < d.b; ++i) {
                const long long span = pvv_pool_span(((volatile int *)ws.tn)[i]);
                if (used + span > ws.cap) {
                    ws.img_flags[i] |= PVV_IMG_OVERFLOW | PVV_IMG_SKIPPED;
                    ws.tn[i] = 0;
                } else {
                    used += span;
                }
            }
            atomicOr(&ws.counters[PVV_CTR_STATUS], (unsigned)PVV_FLAG_POOL_OVERFLOW);
            __threadfence();
        }
        __syncthreads();
    }

    // batch-level scan: pool offsets and voting-chunk ranges.  b is small; chunks of kThreads with a carry.
    long long off = 0;
    int chunk = 0, rank = 0;
    for (int base = 0; base < d.b; base += kThreads) {
        const int i = base + threadIdx.x;
        const int n = i < d.b ? ((volatile int *)ws.tn)[i] : 0;
        // every image starts on a PVV_POOL_ALIGN-pixel boundary of the pool (16-byte aligned bulk copies in K3)
        int total_n;
        const int ex_n = block_excl_scan(pvv_pool_span(n), &total_n);
        const long long my_off = off + ex_n;
        const int c = (n + PVV_CHUNK_PX - 1) / PVV_CHUNK_PX;
        int total_c;
        const int ex_c = block_excl_scan(c, &total_c);
        int total_r = 0, ex_r = 0;
        if (ws.want_rank) ex_r = block_excl_scan(n > 0 ? 1 : 0, &total_r);   // rank of the image among those that vote (k_count_tc2)
        if (i < d.b) {
            ws.img_off[i] = my_off;
            ws.chunk_start[i] = chunk + ex_c;
            ws.img_rank[i] = rank + ex_r;
        }
        off += total_n;
        chunk += total_c;
        rank += total_r;
    }
    if (threadIdx.x == 0) { ws.chunk_start[d.b] = chunk; ws.counters[PVV_CTR_NIMG] = (unsigned)rank; }
    __syncthreads();
    // voting-chunk table: K3 turns a work item into (image, pixel range) with one load instead of a search.
    //   .x image   .y POOL index of the chunk's first pixel   .z pixels | first-chunk-of-its-image << 15 | image rank << 16
    //   .w x | y << 16 of the chunk's middle pixel (written by K1d)
    for (int i = threadIdx.x; i < d.b; i += kThreads) {
        const int n = ws.tn[i], cs = ws.chunk_start[i], rk = ws.img_rank[i];
        const int off0 = (int)ws.img_off[i];
        PVV_CHECK(off0 >= 0 && (long long)off0 + n <= ws.cap && (off0 % PVV_POOL_ALIGN) == 0);
        PVV_CHECK(cs >= 0 && cs + (n + PVV_CHUNK_PX - 1) / PVV_CHUNK_PX <= ws.cap / PVV_CHUNK_PX + d.b + 1);
        for (int p0 = 0, c = 0; p0 < n; p0 += PVV_CHUNK_PX, ++c)
            *reinterpret_cast<int3 *>(&ws.chunk_tab[cs + c]) =
                make_int3(i, off0 + p0, min(PVV_CHUNK_PX, n - p0) | (c == 0 ? 1 << 15 : 0) | (rk << 16));
    }
}

// ---- K1d: scatter (reads the bitmask, not the mask) ------------------------------------------------
#ifndef PVV_TILES_PER_CTA
#define PVV_TILES_PER_CTA 2
#endif
#ifndef PVV_KP_PER_CTA
#define PVV_KP_PER_CTA 64
#endif
constexpr int kTilesPerCta = PVV_TILES_PER_CTA;  // empty tiles (most of an image) cost one load, not one CTA; measured 1/2/3/4/8: 2 is best
constexpr int kKpPerCta = PVV_KP_PER_CTA;     // blockIdx.z = a group of keypoints (a tuning knob: splitting a tile's keypoint planes over
                                 // CTAs measured slower at 1/3/5/6 per CTA than no split, so one CTA takes them all)

__global__ void __launch_bounds__(kThreads) k_scatter(pvv_dims d, const float *vertex, pvv_ws ws)
{
    pvv_pdl_enter();
    const int img = blockIdx.y;
    const int tn = ws.tn[img];
    if (tn == 0) return;
    __shared__ uint32_t s_xy[PVV_TILE_PX];
    __shared__ long long s_off[PVV_TILE_PX];
    for (int tile = blockIdx.x * kTilesPerCta; tile < min(ws.ntiles, (blockIdx.x + 1) * kTilesPerCta); ++tile) {
        const int cnt = ws.tile_cnt[img * ws.ntiles + tile];
        if (cnt == 0) continue;  // CTA-uniform
        // this thread's 4 consecutive pixels of the tile: one nibble of the bitmask
        const uint32_t *bits = ws.fg_bits + ((long long)img * ws.ntiles + tile) * (PVV_TILE_PX / 32);
        const int q0 = threadIdx.x * kPxPerThread;
        const uint32_t flags = (bits[q0 >> 5] >> (q0 & 31)) & 0xFu;
        const int p0 = tile * PVV_TILE_PX + q0;
        int total;
        int rank = block_excl_scan(__popc(flags), &total);  // total == cnt: the bitmask is what was counted
        if (flags) {
            int y = p0 / d.w, x = p0 - y * d.w;
#pragma unroll
            for (int k = 0; k < kPxPerThread; ++k) {
                if ((flags >> k) & 1u) {
                    s_xy[rank] = (uint32_t)x | ((uint32_t)y << 16);
                    s_off[rank] = img * d.vs[0] + y * d.vs[1] + x * d.vs[2];
                    ++rank;
                }
                if (++x == d.w) { x = 0; ++y; }
            }
        }
        __syncthreads();
        const int toff = ws.tile_off[img * ws.ntiles + tile];
        const long long base = ws.img_off[img] + toff;
        const int cs = ws.chunk_start[img];
        // consecutive threads take consecutive pixels (coalesced along x inside a keypoint plane); every thread walks the
        // keypoints of its pixel
        const int vi0 = blockIdx.z * kKpPerCta, vi1 = min(d.vn, vi0 + kKpPerCta);
        for (int j = threadIdx.x; j < total; j += kThreads) {
            PVV_CHECK(base + j >= 0 && base + j < ws.cap);
            if (blockIdx.z == 0) {
                ws.pix_xy[base + j] = s_xy[j];
                // the middle pixel of a voting chunk is the origin of that chunk's local frame in K3
                const int t = toff + j, c = t / PVV_CHUNK_PX;
                if (t - c * PVV_CHUNK_PX == (min(PVV_CHUNK_PX, tn - c * PVV_CHUNK_PX) >> 1)) {
                    ws.chunk_tab[cs + c].w = (int)s_xy[j];
                    // the middle pixel of the image's middle chunk (its median pixel in raster order) is the origin of
                    // the image's frame in the tensor-pipe K3
                    if (c == ((tn + PVV_CHUNK_PX - 1) / PVV_CHUNK_PX) >> 1) ws.img_org[img] = s_xy[j];
                }
            }
            const float *src = vertex + s_off[j];
            float2 *dst = ws.pix_dir + base + j;
            // 3 keypoints (6 loads) in flight per thread: measured best -- 18 in flight is 1.7x slower (the planes of one
            // image are 1.2 MB apart: too many open DRAM pages / TLB entries per warp)
#pragma unroll 3
            for (int vi = vi0; vi < vi1; ++vi) {
                float2 v;
                v.x = __ldg(src + vi * d.vs[3]);
                v.y = __ldg(src + vi * d.vs[3] + d.vs[4]);
                dst[(long long)vi * ws.cap] = v;
            }
        }
        __syncthreads();  // s_xy / s_off are reused by the next tile
    }
}

}  // namespace

cudaError_t pvv_launch_compact(const pvv_dims &d, const void *mask, const float *vertex, const float *selection,
                               const pvv_ws &ws, cudaStream_t st, int *launches)
{
    const dim3 grid(ws.ntiles, d.b);
    switch (d.mask_dtype) {
        case PVV_MASK_U8: launch_tile_count<PVV_MASK_U8>(d, mask, ws, grid, st); break;
        case PVV_MASK_I32: launch_tile_count<PVV_MASK_I32>(d, mask, ws, grid, st); break;
        case PVV_MASK_I64: launch_tile_count<PVV_MASK_I64>(d, mask, ws, grid, st); break;
        case PVV_MASK_LOGITS_F32: launch_tile_count<PVV_MASK_LOGITS_F32>(d, mask, ws, grid, st); break;
        default: launch_tile_count<PVV_MASK_F32>(d, mask, ws, grid, st); break;
    }
    ++*launches;
    if ((long long)d.max_num < (long long)d.h * d.w * (d.fg_rule == PVV_FG_BYTE ? 255 : 1)) {
        const dim3 rgrid((ws.ntiles + kThreads / 32 - 1) / (kThreads / 32), d.b);
        pvv_launch_pdl(k_tile_recount, rgrid, dim3(kThreads), st, d, selection, ws);
        ++*launches;
    }
    pvv_launch_pdl(k_scan, dim3(d.b), dim3(kThreads), st, d, ws);
    pvv_launch_pdl(k_scatter, dim3((ws.ntiles + kTilesPerCta - 1) / kTilesPerCta, d.b, (d.vn + kKpPerCta - 1) / kKpPerCta),
                   dim3(kThreads), st, d, vertex, ws);
    *launches += 2;
    return cudaGetLastError();
}
