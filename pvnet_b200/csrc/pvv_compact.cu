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
// Four launches: tile counts -> (optional) recount under subsampling -> per-image scan + batch scan by the
// last block -> scatter.  The mask is read through its real strides and dtype; the vertex tensor is read
// through its real strides (the network emits a permuted view), only at foreground pixels.
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

// foreground flags of this thread's 4 consecutive pixels (bit k = pixel p0+k), plus the byte-value sum
__device__ __forceinline__ uint32_t load_flags(const pvv_dims &d, const void *mask, int img, int p0, uint32_t *bsum)
{
    const int npx = d.h * d.w;
    uint32_t flags = 0, s = 0;
    int y = p0 / d.w, x = p0 - y * d.w;
#pragma unroll
    for (int k = 0; k < kPxPerThread; ++k) {
        if (p0 + k < npx) {
            const uint32_t v = pvv_mask_value(mask, d.mask_dtype, img * d.ms[0] + y * d.ms[1] + x * d.ms[2], d.fg_rule);
            s += v;
            flags |= (v != 0u) << k;
        }
        if (++x == d.w) { x = 0; ++y; }
    }
    *bsum = s;
    return flags;
}

// ---- K1a: per-tile foreground count + per-image foreground sum ------------------------------------
__global__ void __launch_bounds__(kThreads) k_tile_count(pvv_dims d, const void *mask, pvv_ws ws)
{
    const int img = blockIdx.y, tile = blockIdx.x;
    uint32_t bsum;
    const uint32_t flags = load_flags(d, mask, img, tile * PVV_TILE_PX + threadIdx.x * kPxPerThread, &bsum);
    int n = __popc(flags);
    int s = (int)bsum;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        n += __shfl_xor_sync(0xffffffffu, n, o);
        s += __shfl_xor_sync(0xffffffffu, s, o);
    }
    __shared__ int s_n[kThreads / 32], s_s[kThreads / 32];
    if ((threadIdx.x & 31) == 0) { s_n[threadIdx.x >> 5] = n; s_s[threadIdx.x >> 5] = s; }
    __syncthreads();
    if (threadIdx.x == 0) {
        int tn = 0, ts = 0;
#pragma unroll
        for (int i = 0; i < kThreads / 32; ++i) { tn += s_n[i]; ts += s_s[i]; }
        ws.tile_cnt[img * ws.ntiles + tile] = tn;
        if (ts) atomicAdd((unsigned long long *)&ws.img_fg[img], (unsigned long long)ts);
    }
}

// ---- K1a': recount tiles of images that get subsampled ---------------------------------------------
__global__ void __launch_bounds__(kThreads) k_tile_recount(pvv_dims d, const void *mask, const float *selection,
                                                           pvv_ws ws)
{
    const int img = blockIdx.y, tile = blockIdx.x;
    const long long fg = ws.img_fg[img];
    if (fg <= d.max_num || fg < d.min_num) return;
    const float prob = pvv_subsample_prob(d.max_num, fg);
    const int p0 = tile * PVV_TILE_PX + threadIdx.x * kPxPerThread;
    uint32_t bsum;
    uint32_t flags = load_flags(d, mask, img, p0, &bsum);
    int n = 0;
#pragma unroll
    for (int k = 0; k < kPxPerThread; ++k)
        if ((flags >> k) & 1u) n += sel_value(d, selection, img, (uint32_t)(p0 + k)) < prob;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
    __shared__ int s_n[kThreads / 32];
    if ((threadIdx.x & 31) == 0) s_n[threadIdx.x >> 5] = n;
    __syncthreads();
    if (threadIdx.x == 0) {
        int tn = 0;
#pragma unroll
        for (int i = 0; i < kThreads / 32; ++i) tn += s_n[i];
        ws.tile_cnt[img * ws.ntiles + tile] = tn;
    }
}

// ---- K1c: per-image exclusive scan of tile counts; the last block scans the batch ------------------
__global__ void __launch_bounds__(kThreads) k_scan(pvv_dims d, pvv_ws ws)
{
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
        __threadfence();
        s_last = atomicAdd(&ws.counters[PVV_CTR_TICKET], 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();

    // batch-level scan: pool offsets and voting-chunk ranges.  b is small; chunks of kThreads with a carry.
    long long off = 0;
    int chunk = 0;
    for (int base = 0; base < d.b; base += kThreads) {
        const int i = base + threadIdx.x;
        int n = i < d.b ? ((volatile int *)ws.tn)[i] : 0;
        // prefix of n (pixels) -- int is enough per 256 images of <= 2^31/256... keep it safe with two scans
        int total_n;
        const int ex_n = block_excl_scan(n, &total_n);
        long long my_off = off + ex_n;
        if (i < d.b && my_off + n > ws.cap) {  // pool overflow: drop the image, flag it
            atomicOr(&ws.counters[PVV_CTR_STATUS], (unsigned)PVV_FLAG_POOL_OVERFLOW);
            ws.img_flags[i] |= PVV_IMG_OVERFLOW | PVV_IMG_SKIPPED;
            ws.tn[i] = 0;
            n = 0;
            my_off = my_off > ws.cap ? ws.cap : my_off;
        }
        const int c = (n + PVV_CHUNK_PX - 1) / PVV_CHUNK_PX;
        int total_c;
        const int ex_c = block_excl_scan(c, &total_c);
        if (i < d.b) {
            ws.img_off[i] = my_off;
            ws.chunk_start[i] = chunk + ex_c;
        }
        off += total_n;
        chunk += total_c;
    }
    if (threadIdx.x == 0) ws.chunk_start[d.b] = chunk;
}

// ---- K1d: scatter ----------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) k_scatter(pvv_dims d, const void *mask, const float *vertex,
                                                      const float *selection, pvv_ws ws)
{
    const int img = blockIdx.y, tile = blockIdx.x;
    const int tn = ws.tn[img];
    if (tn == 0) return;
    const int cnt = ws.tile_cnt[img * ws.ntiles + tile];
    if (cnt == 0) return;
    const int iflags = ws.img_flags[img];
    const int p0 = tile * PVV_TILE_PX + threadIdx.x * kPxPerThread;
    uint32_t bsum;
    uint32_t flags = load_flags(d, mask, img, p0, &bsum);
    if (iflags & PVV_IMG_SUBSAMPLED) {
        const float prob = pvv_subsample_prob(d.max_num, ws.img_fg[img]);
#pragma unroll
        for (int k = 0; k < kPxPerThread; ++k)
            if (((flags >> k) & 1u) && !(sel_value(d, selection, img, (uint32_t)(p0 + k)) < prob)) flags &= ~(1u << k);
    }
    __shared__ int s_list[PVV_TILE_PX];
    int total;
    int rank = block_excl_scan(__popc(flags), &total);
#pragma unroll
    for (int k = 0; k < kPxPerThread; ++k)
        if ((flags >> k) & 1u) s_list[rank++] = p0 + k;
    __syncthreads();
    // total == cnt by construction (same deterministic predicate as the count pass)
    const long long base = ws.img_off[img] + ws.tile_off[img * ws.ntiles + tile];
    for (int j = threadIdx.x; j < total; j += kThreads) {
        const int p = s_list[j];
        const int y = p / d.w, x = p - y * d.w;
        ws.pix_xy[base + j] = (uint32_t)x | ((uint32_t)y << 16);
    }
    const int work = total * d.vn;
    for (int e = threadIdx.x; e < work; e += kThreads) {
        const int vi = e / total, j = e - vi * total;
        const int p = s_list[j];
        const int y = p / d.w, x = p - y * d.w;
        const long long vo = img * d.vs[0] + y * d.vs[1] + x * d.vs[2] + vi * d.vs[3];
        float2 v;
        v.x = __ldg(vertex + vo);
        v.y = __ldg(vertex + vo + d.vs[4]);
        ws.pix_dir[(long long)vi * ws.cap + base + j] = v;
    }
}

}  // namespace

cudaError_t pvv_launch_compact(const pvv_dims &d, const void *mask, const float *vertex, const float *selection,
                               const pvv_ws &ws, cudaStream_t st, int *launches)
{
    const dim3 grid(ws.ntiles, d.b);
    k_tile_count<<<grid, kThreads, 0, st>>>(d, mask, ws);
    ++*launches;
    if ((long long)d.max_num < (long long)d.h * d.w * (d.fg_rule == PVV_FG_BYTE ? 255 : 1)) {
        k_tile_recount<<<grid, kThreads, 0, st>>>(d, mask, selection, ws);
        ++*launches;
    }
    k_scan<<<d.b, kThreads, 0, st>>>(d, ws);
    k_scatter<<<grid, kThreads, 0, st>>>(d, mask, vertex, selection, ws);
    *launches += 2;
    return cudaGetLastError();
}
