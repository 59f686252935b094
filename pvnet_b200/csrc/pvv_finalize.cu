// pvv_finalize.cu -- K4 (argmax + winner re-vote + least-squares refinement) and K5 (weighted covariance).
//
// K4 replaces, per (image, keypoint) and with no host round-trip,
//   torch.max(counts, 0) / gather / keep-best                      ransac_voting_gpu.py:568-575
//   winner re-vote                                                 :585-590
//   weighted normal equations + 2x2 solve (b_inv)                  :593-601, :503-518
// K5 replaces the covariance algebra of estimate_voting_distribution_with_mean, :375, :394-401.
// Integer decisions (argmax, inlier flags) are bit-exact; the sums are accumulated in float64 from the same
// float32 terms the reference forms (its own float32 sum order is library-defined; parity bar 1e-4 relative).
#include "pvv_internal.cuh"

namespace {

#ifndef PVV_FIN_THREADS
#define PVV_FIN_THREADS 256
#endif
constexpr int kFinThreads = PVV_FIN_THREADS;

__device__ __forceinline__ double block_sum(double v, double *s_buf)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) s_buf[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0.0;
#pragma unroll
    for (int i = 0; i < kFinThreads / 32; ++i) t += s_buf[i];
    return t;
}

// five sums at once (the normal equations): one pair of barriers instead of five
__device__ __forceinline__ void block_sum5(double (&v)[5], double *s_buf5 /* [5 * kFinThreads / 32] */)
{
#pragma unroll
    for (int i = 0; i < 5; ++i)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v[i] += __shfl_xor_sync(0xffffffffu, v[i], o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0)
#pragma unroll
        for (int i = 0; i < 5; ++i) s_buf5[i * (kFinThreads / 32) + (threadIdx.x >> 5)] = v[i];
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 5; ++i) {
        double t = 0.0;
#pragma unroll
        for (int w = 0; w < kFinThreads / 32; ++w) t += s_buf5[i * (kFinThreads / 32) + w];
        v[i] = t;
    }
}

// max count with lowest-index tie rule (torch.max(dim) returns the first maximal index)
__device__ __forceinline__ void block_argmax(const int *row, int hn, int *best_cnt, int *best_h)
{
    __shared__ long long s_key[kFinThreads / 32];
    // key = count in the high word, (INT_MAX - h) in the low word: max key <=> max count, then min h
    long long key = -1;
    for (int h = threadIdx.x; h < hn; h += kFinThreads) {
        const long long k = ((long long)row[h] << 32) | (unsigned)(0x7fffffff - h);
        key = k > key ? k : key;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const long long other = __shfl_xor_sync(0xffffffffu, key, o);
        key = other > key ? other : key;
    }
    if ((threadIdx.x & 31) == 0) s_key[threadIdx.x >> 5] = key;
    __syncthreads();
    long long best = -1;
#pragma unroll
    for (int i = 0; i < kFinThreads / 32; ++i) best = s_key[i] > best ? s_key[i] : best;
    *best_cnt = (int)(best >> 32);
    *best_h = 0x7fffffff - (int)(best & 0xffffffffLL);
    __syncthreads();
}

struct v3_dbg {
    int32_t *win_idx, *win_count;
    float *win_pts;
    uint8_t *win_inlier;
    double *normal_eq;
    long long dbg_cap;
};

// grid (vn, b, PVV_FIN_SPLIT): the pixels of one (image, keypoint) are shared by PVV_FIN_SPLIT CTAs; each writes its
// partial normal equations to its own slot and the last one to arrive adds the slots in a fixed order (deterministic,
// unlike float atomics) and solves.
__global__ void __launch_bounds__(kFinThreads) k_v3_finalize(pvv_dims d, pvv_ws ws, float *out_kp, float *out_conf, v3_dbg g)
{
    __shared__ double s_buf[kFinThreads / 32];
    __shared__ double s_buf5[5 * kFinThreads / 32];
    pvv_pdl_enter();
    __shared__ float2 s_pt;
    __shared__ bool s_last;
    const int img = blockIdx.y, vi = blockIdx.x, part = blockIdx.z;
    const int tn = ws.tn[img];
    const long long o = (long long)img * d.vn + vi;
    if (tn == 0) {  // skipped image: zeros (:537-541)
        if (threadIdx.x == 0 && part == 0) {
            out_kp[o * 2] = 0.f; out_kp[o * 2 + 1] = 0.f;
            if (out_conf) out_conf[o] = 0.f;                                          // v5: pts_conf zeros (:787-789)
            if (g.win_idx) g.win_idx[o] = 0;
            if (g.win_count) g.win_count[o] = 0;
            if (g.win_pts) { g.win_pts[o * 2] = 0.f; g.win_pts[o * 2 + 1] = 0.f; }
            if (g.normal_eq) for (int i = 0; i < 5; ++i) g.normal_eq[o * 5 + i] = 0.0;
        }
        return;
    }
    const long long row = o * ws.hn_pad;
    int best_cnt, best_h;
    block_argmax(ws.counts + row, d.hn, &best_cnt, &best_h);
    // all_win_ratio starts at 0 and is only replaced by a strictly larger ratio (:573): count 0 keeps (0,0)
    const float2 win = best_cnt > 0 ? ws.hyp[row + best_h] : make_float2(0.f, 0.f);

    const long long off = ws.img_off[img];
    double acc[5] = {0, 0, 0, 0, 0};  // ATA00, ATA01, ATA11, ATb0, ATb1
    const float2 *pdir = ws.pix_dir + (long long)vi * ws.cap + off;
    constexpr int STEP = kFinThreads * PVV_FIN_SPLIT;
    for (int t0 = part * kFinThreads + threadIdx.x; t0 < tn; t0 += 2 * STEP) {
        // two pixels per step, loads first: the loop is a chain of L2 round trips otherwise
        const int t1 = t0 + STEP;
        const bool has1 = t1 < tn;
        const uint32_t xy0 = ws.pix_xy[off + t0], xy1 = has1 ? ws.pix_xy[off + t1] : 0u;
        const float2 dir0 = pdir[t0], dir1 = has1 ? pdir[t1] : make_float2(0.f, 0.f);
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            if (u == 1 && !has1) break;
            const int t = u ? t1 : t0;
            const uint32_t xy = u ? xy1 : xy0;
            const float2 dir = u ? dir1 : dir0;
            const float cx = (float)(xy & 0xffffu), cy = (float)(xy >> 16);
            const bool inl = pvv_exact_test(dir.x, dir.y, pvv_norm1(dir.x, dir.y), cx, cy, win.x, win.y, d.thresh);
            if (g.win_inlier) g.win_inlier[(long long)vi * g.dbg_cap + off + t] = inl ? 1 : 0;
            if (inl) {
                const float nx = dir.y, ny = -dir.x;                                 // :586-587
                const float bb = __fadd_rn(__fmul_rn(nx, cx), __fmul_rn(ny, cy));    // :596 (fp32 products, 2-term sum)
                acc[0] += (double)nx * nx; acc[1] += (double)nx * ny; acc[2] += (double)ny * ny;
                acc[3] += (double)__fmul_rn(nx, bb);                                 // :598
                acc[4] += (double)__fmul_rn(ny, bb);
            }
        }
    }
    block_sum5(acc, s_buf5);
    double s00 = acc[0], s01 = acc[1], s11 = acc[2], b0 = acc[3], b1 = acc[4];
    if (threadIdx.x == 0) {
        double *slot = ws.fin_part + (o * PVV_FIN_SPLIT + part) * 5;
        slot[0] = s00; slot[1] = s01; slot[2] = s11; slot[3] = b0; slot[4] = b1;
        __threadfence();
        s_last = atomicAdd(&ws.fin_ticket[o], 1u) == PVV_FIN_SPLIT - 1;
    }
    __syncthreads();
    if (!s_last) return;
    if (threadIdx.x == 0) {
        __threadfence();
        s00 = s01 = s11 = b0 = b1 = 0.0;
        const volatile double *slots = ws.fin_part + o * PVV_FIN_SPLIT * 5;
        for (int p = 0; p < PVV_FIN_SPLIT; ++p) {  // fixed order: bit-reproducible sums
            s00 += slots[p * 5 + 0]; s01 += slots[p * 5 + 1]; s11 += slots[p * 5 + 2];
            b0 += slots[p * 5 + 3]; b1 += slots[p * 5 + 4];
        }
        const double det = s00 * s11 - s01 * s01;
        double x = b0, y = b1;  // b_inv's identity fallback (:516-517), applied per keypoint
        if (det != 0.0 && isfinite(det)) {
            x = (s11 * b0 - s01 * b1) / det;
            y = (s00 * b1 - s01 * b0) / det;
        }
        out_kp[o * 2] = (float)x;
        out_kp[o * 2 + 1] = (float)y;
        s_pt = make_float2((float)x, (float)y);
        if (g.win_idx) g.win_idx[o] = best_h;
        if (g.win_count) g.win_count[o] = best_cnt;
        if (g.win_pts) { g.win_pts[o * 2] = win.x; g.win_pts[o * 2 + 1] = win.y; }
        if (g.normal_eq) {
            double *q = g.normal_eq + o * 5;
            q[0] = s00; q[1] = s01; q[2] = s11; q[3] = b0; q[4] = b1;
        }
    }
    if (!out_conf) return;
    // v5 confidence: re-vote the refined keypoint at the fixed threshold 0.999, conf = count / tn      (:854-856)
    __syncthreads();
    const float2 pt = s_pt;
    double c = 0.0;
    for (int t = threadIdx.x; t < tn; t += kFinThreads) {
        const uint32_t xy = ws.pix_xy[off + t];
        const float2 dir = ws.pix_dir[(long long)vi * ws.cap + off + t];
        c += pvv_exact_test(dir.x, dir.y, pvv_norm1(dir.x, dir.y), (float)(xy & 0xffffu), (float)(xy >> 16), pt.x, pt.y,
                            0.999f) ? 1.0 : 0.0;
    }
    c = block_sum(c, s_buf);
    if (threadIdx.x == 0) out_conf[o] = __fdiv_rn((float)c, (float)tn);
}

__global__ void __launch_bounds__(kFinThreads) k_cov_finalize(pvv_dims d, pvv_ws ws, const float *mean, float *out_cov)
{
    __shared__ double s_buf[kFinThreads / 32];
    pvv_pdl_enter();
    const int img = blockIdx.y, vi = blockIdx.x;
    const int tn = ws.tn[img];
    const long long o = (long long)img * d.vn + vi;
    const float mx = mean[o * 2], my = mean[o * 2 + 1];
    const long long row = o * ws.hn_pad;
    double c00 = 0, c01 = 0, c10 = 0, c11 = 0, sw = 0;
    if (tn == 0) {
        // skipped image: hypotheses are zeros and every ratio is one (:343-348) -> threshold 0.9 keeps all
        if (threadIdx.x == 0) {
            const float dx = __fsub_rn(0.f, mx), dy = __fsub_rn(0.f, my);
            c00 = (double)dx * dx * d.hn; c01 = (double)dx * dy * d.hn; c10 = c01; c11 = (double)dy * dy * d.hn;
            sw = (double)d.hn;
        }
    } else {
        int best_cnt, best_h;
        block_argmax(ws.counts + row, d.hn, &best_cnt, &best_h);
        const float ftn = (float)tn;
        const float thr = __fsub_rn(__fdiv_rn((float)best_cnt, ftn), 0.1f);          // :394
        for (int h = threadIdx.x; h < d.hn; h += kFinThreads) {
            const float r = __fdiv_rn((float)ws.counts[row + h], ftn);              // :375
            const float w = r < thr ? 0.f : r;                                      // :395
            const float2 hv = ws.hyp[row + h];
            const float dx = __fsub_rn(hv.x, mx), dy = __fsub_rn(hv.y, my);          // :398
            const float wx = __fmul_rn(dx, w), wy = __fmul_rn(dy, w);                // :399
            c00 += (double)dx * wx; c01 += (double)dx * wy; c10 += (double)dy * wx; c11 += (double)dy * wy;  // :400
            sw += (double)w;
        }
    }
    c00 = block_sum(c00, s_buf); c01 = block_sum(c01, s_buf); c10 = block_sum(c10, s_buf);
    c11 = block_sum(c11, s_buf); sw = block_sum(sw, s_buf);
    if (threadIdx.x == 0) {
        const double den = sw + 1e-3;                                               // :401
        out_cov[o * 4 + 0] = (float)(c00 / den);
        out_cov[o * 4 + 1] = (float)(c01 / den);
        out_cov[o * 4 + 2] = (float)(c10 / den);
        out_cov[o * 4 + 3] = (float)(c11 / den);
    }
}

// ransac_motion_voting (:966-987): mean over the foreground pixels of (vertex + coordinate), fp64 accumulation of the
// fp32 terms torch forms (cur_vert[cur_mask] + coords, :983)
__global__ void __launch_bounds__(kFinThreads) k_motion_mean(pvv_dims d, pvv_ws ws, float *out_kp)
{
    __shared__ double s_buf[kFinThreads / 32];
    pvv_pdl_enter();
    const int img = blockIdx.y, vi = blockIdx.x;
    const int tn = ws.tn[img];
    const long long o = (long long)img * d.vn + vi, off = ws.img_off[img];
    double sx = 0, sy = 0;
    for (int t = threadIdx.x; t < tn; t += kFinThreads) {
        const uint32_t xy = ws.pix_xy[off + t];
        const float2 dir = ws.pix_dir[(long long)vi * ws.cap + off + t];
        sx += (double)__fadd_rn(dir.x, (float)(xy & 0xffffu));
        sy += (double)__fadd_rn(dir.y, (float)(xy >> 16));
    }
    sx = block_sum(sx, s_buf);
    sy = block_sum(sy, s_buf);
    if (threadIdx.x == 0) {
        out_kp[o * 2] = tn > 0 ? (float)(sx / tn) : 0.f;      // no foreground: zeros (:978-980)
        out_kp[o * 2 + 1] = tn > 0 ? (float)(sy / tn) : 0.f;
    }
}

// counts [b][vn][hn_pad] -> reference layout [b][hn][vn]
__global__ void k_export_counts(pvv_dims d, pvv_ws ws, int32_t *counts_out)
{
    pvv_pdl_enter();
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long total = (long long)d.b * d.hn * d.vn;
    if (e >= total) return;
    const int vi = (int)(e % d.vn);
    const long long r = e / d.vn;
    const int hi = (int)(r % d.hn);
    const int img = (int)(r / d.hn);
    PVV_CHECK(img < d.b && vi < d.vn && hi < d.hn);
    counts_out[e] = ws.counts[((long long)img * d.vn + vi) * ws.hn_pad + hi];
}

}  // namespace

cudaError_t pvv_launch_v3_finalize(const pvv_dims &d, const pvv_ws &ws, float *out_kp, float *out_conf,
                                   const pvv_debug *dbg, cudaStream_t st, int *launches)
{
    v3_dbg g = {nullptr, nullptr, nullptr, nullptr, nullptr, 0};
    if (dbg) {
        g.win_idx = dbg->win_idx; g.win_count = dbg->win_count; g.win_pts = dbg->win_pts;
        g.win_inlier = dbg->win_inlier; g.normal_eq = dbg->normal_eq; g.dbg_cap = dbg->pool_capacity;
    }
    pvv_launch_pdl(k_v3_finalize, dim3(d.vn, d.b, PVV_FIN_SPLIT), dim3(kFinThreads), st, d, ws, out_kp, out_conf, g);
    ++*launches;
    return cudaGetLastError();
}

cudaError_t pvv_launch_motion_mean(const pvv_dims &d, const pvv_ws &ws, float *out_kp, cudaStream_t st, int *launches)
{
    pvv_launch_pdl(k_motion_mean, dim3(d.vn, d.b), dim3(kFinThreads), st, d, ws, out_kp);
    ++*launches;
    return cudaGetLastError();
}

cudaError_t pvv_launch_cov_finalize(const pvv_dims &d, const pvv_ws &ws, const float *mean, float *out_cov,
                                    cudaStream_t st, int *launches)
{
    pvv_launch_pdl(k_cov_finalize, dim3(d.vn, d.b), dim3(kFinThreads), st, d, ws, mean, out_cov);
    ++*launches;
    return cudaGetLastError();
}

cudaError_t pvv_launch_export_counts(const pvv_dims &d, const pvv_ws &ws, int32_t *counts_out, cudaStream_t st,
                                     int *launches)
{
    const long long total = (long long)d.b * d.hn * d.vn;
    pvv_launch_pdl(k_export_counts, dim3((unsigned)((total + 255) / 256)), dim3(256), st, d, ws, counts_out);
    ++*launches;
    return cudaGetLastError();
}

cudaError_t pvv_launch_debug_export(const pvv_dims &d, const pvv_ws &ws, const pvv_debug *dbg, cudaStream_t st,
                                    int *launches)
{
    if (!dbg) return cudaSuccess;
    cudaError_t e = cudaSuccess;
    if (dbg->tn) e = cudaMemcpyAsync(dbg->tn, ws.tn, sizeof(int) * d.b, cudaMemcpyDeviceToDevice, st);
    if (!e && dbg->img_off) e = cudaMemcpyAsync(dbg->img_off, ws.img_off, sizeof(long long) * d.b, cudaMemcpyDeviceToDevice, st);
    if (!e && dbg->pix_xy) e = cudaMemcpyAsync(dbg->pix_xy, ws.pix_xy, sizeof(uint32_t) * ws.cap, cudaMemcpyDeviceToDevice, st);
    if (!e && dbg->pix_dir) e = cudaMemcpyAsync(dbg->pix_dir, ws.pix_dir, sizeof(float2) * ws.cap * d.vn, cudaMemcpyDeviceToDevice, st);
    if (!e && dbg->status) e = cudaMemcpyAsync(dbg->status, &ws.counters[PVV_CTR_STATUS], sizeof(int), cudaMemcpyDeviceToDevice, st);
    if (!e && dbg->amb_count) e = cudaMemcpyAsync(dbg->amb_count, &ws.counters[PVV_CTR_AMB_LO], sizeof(uint64_t), cudaMemcpyDeviceToDevice, st);
    if (!e && dbg->counts) {
        const long long total = (long long)d.b * d.hn * d.vn;
        pvv_launch_pdl(k_export_counts, dim3((unsigned)((total + 255) / 256)), dim3(256), st, d, ws, dbg->counts);
        ++*launches;
        e = cudaGetLastError();
    }
    return e;
}
