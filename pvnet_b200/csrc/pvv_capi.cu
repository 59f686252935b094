// pvv_capi.cu -- the extern "C" surface of libpvvote.so (include/pvvote.h): argument validation, workspace
// carving, launch sequencing on the caller's stream.  No torch types, no global mutable state, no exit().
#include <math.h>
#include <string.h>

#include "pvv_internal.cuh"

// Scratch of pvv_v3_noalloc: a file-scope device array (no cudaMalloc anywhere on that path).  In host code this name is a
// host-side stand-in, NOT a device address: its device address is taken once per device with cudaGetSymbolAddress.
#define PVV_STATIC_WS_BYTES (32u << 20)
static __device__ __align__(256) unsigned char g_pvv_static_ws[PVV_STATIC_WS_BYTES];

namespace {

thread_local int t_last_launches = 0;
thread_local int t_last_cuda_error = 0;

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

int fail_cuda(cudaError_t e)
{
    t_last_cuda_error = (int)e;
    return PVV_ERR_CUDA;
}

// pvv_job.status: the call's PVV_FLAG_* bits, copied on the job's stream (no host synchronisation)
cudaError_t export_status(const pvv_job *j, const pvv_ws &ws)
{
    if (!j->status) return cudaSuccess;
    return cudaMemcpyAsync(j->status, &ws.counters[PVV_CTR_STATUS], sizeof(int32_t), cudaMemcpyDeviceToDevice,
                           (cudaStream_t)j->stream);
}

long long default_cap(const pvv_job *j)
{
    const long long npx = (long long)j->h * j->w;
    long long per = npx;
    if (j->max_num >= 0 && j->max_num < npx) {
        // fg <= max_num is kept whole; above it a Bernoulli(max_num/fg) draw keeps ~max_num +- sqrt(max_num)
        const long long bound = (long long)j->max_num + (long long)(8.0 * sqrt((double)j->max_num)) + 256;
        per = bound < npx ? bound : npx;
    }
    return per * j->b;
}

// pool pixels allocated for a requested capacity: every image may lose up to PVV_POOL_ALIGN - 1 pixels to the alignment
// of its start, and the bulk copies of K3 read whole 16-byte units past a chunk's last pixel
long long pool_alloc(const pvv_job *j)
{
    const long long want = j->pool_capacity > 0 ? j->pool_capacity : default_cap(j);
    return (want + (long long)PVV_POOL_ALIGN * (j->b + 1) + 63) / 64 * 64;
}

int validate(const pvv_job *j)
{
    if (!j || !j->mask || !j->vertex || !j->workspace) return PVV_ERR_INVALID;
    if (j->b <= 0 || j->h <= 0 || j->w <= 0 || j->vn <= 0 || j->hn_total <= 0) return PVV_ERR_INVALID;
    if (j->mask_dtype < PVV_MASK_U8 || j->mask_dtype > PVV_MASK_LOGITS_F32) return PVV_ERR_INVALID;
    if (j->mask_dtype == PVV_MASK_LOGITS_F32 && j->seg_classes < 1) return PVV_ERR_INVALID;
    if (j->fg_rule != PVV_FG_BYTE && j->fg_rule != PVV_FG_EQ1) return PVV_ERR_INVALID;
    if (!(j->inlier_thresh == j->inlier_thresh)) return PVV_ERR_INVALID;
    if (j->vn > PVV_MAX_VN || j->h > 65535 || j->w > 65535 || (long long)j->h * j->w > (1LL << 23) || j->b > 65535)
        return PVV_ERR_UNSUPPORTED;
    if (pool_alloc(j) >= (1LL << 31)) return PVV_ERR_UNSUPPORTED;  // pool pixel indices are 32-bit
    return PVV_OK;
}

struct layout {
    size_t zero_bytes, total_bytes;
    size_t o_counters, o_img_fg, o_counts, o_ticket, o_part, o_flags, o_tn, o_img_off, o_chunk, o_ctab, o_tile_cnt, o_tile_off, o_bits, o_xy, o_dir, o_pq, o_hyp;
    size_t o_rank, o_org, o_tcq, o_tcq_cnt;
    long long cap;
    int ntiles, hn_pad, tcq_slice_cap;
};

layout make_layout(const pvv_job *j)
{
    layout L;
    L.cap = pool_alloc(j);
    L.ntiles = (int)(((long long)j->h * j->w + PVV_TILE_PX - 1) / PVV_TILE_PX);
    L.hn_pad = (j->hn_total + 3) / 4 * 4;
    size_t o = 0;
    auto take = [&](size_t bytes) { const size_t at = o; o = align_up(o + bytes, 256); return at; };
    L.o_counters = take(16 * sizeof(unsigned));
    L.o_img_fg = take(sizeof(long long) * j->b);
    L.o_counts = take(sizeof(int) * (size_t)j->b * j->vn * L.hn_pad);
    L.o_ticket = take(sizeof(unsigned) * (size_t)j->b * j->vn);
    L.o_tcq_cnt = take(sizeof(int) * PVV_TCQ_SLICES);
    L.zero_bytes = o;
    L.o_flags = take(sizeof(int) * j->b);
    L.o_tn = take(sizeof(int) * j->b);
    L.o_img_off = take(sizeof(long long) * j->b);
    L.o_chunk = take(sizeof(int) * (j->b + 1));
    L.o_ctab = take(sizeof(int4) * ((size_t)L.cap / PVV_CHUNK_PX + (size_t)j->b + 1));
    L.o_tile_cnt = take(sizeof(int) * (size_t)j->b * L.ntiles);
    L.o_tile_off = take(sizeof(int) * (size_t)j->b * L.ntiles);
    L.o_bits = take(sizeof(uint32_t) * (size_t)j->b * L.ntiles * (PVV_TILE_PX / 32));
    L.o_xy = take(sizeof(uint32_t) * (size_t)L.cap);
    L.o_dir = take(sizeof(float2) * (size_t)L.cap * j->vn);
    L.o_hyp = take(sizeof(float2) * (size_t)j->b * j->vn * L.hn_pad);
    L.o_pq = take(sizeof(float4) * (size_t)L.cap * j->vn);
    L.o_part = take(sizeof(double) * (size_t)j->b * j->vn * PVV_FIN_SPLIT * 5);
    L.o_rank = take(sizeof(int) * j->b);
    L.o_org = take(sizeof(uint32_t) * j->b);
    {   // queue of the tensor-pipe K3's flagged (hypothesis, 16-pixel group) pairs: 64 per unit of work on average (measured
        // 57 at thresh 0.999, 8 at 0.99; a unit has 8192), within [2^17, 2^22] entries; whatever does not fit is resolved in
        // line by the kernel
        const long long units = ((long long)L.cap / PVV_CHUNK_PX + j->b) * j->vn * ((j->hn_total + 511) / 512);
        long long entries = 64 * units;
        entries = entries < (1LL << 17) ? (1LL << 17) : entries > (1LL << 22) ? (1LL << 22) : entries;
        L.tcq_slice_cap = (int)(entries / PVV_TCQ_SLICES);
        L.o_tcq = take(sizeof(uint4) * (size_t)L.tcq_slice_cap * PVV_TCQ_SLICES);
    }
    L.total_bytes = o;
    return L;
}

pvv_ws carve(const layout &L, void *base)
{
    char *p = (char *)base;
    pvv_ws ws;
    ws.counters = (unsigned *)(p + L.o_counters);
    ws.img_fg = (long long *)(p + L.o_img_fg);
    ws.counts = (int *)(p + L.o_counts);
    ws.img_flags = (int *)(p + L.o_flags);
    ws.tn = (int *)(p + L.o_tn);
    ws.img_off = (long long *)(p + L.o_img_off);
    ws.chunk_start = (int *)(p + L.o_chunk);
    ws.chunk_tab = (int4 *)(p + L.o_ctab);
    ws.tile_cnt = (int *)(p + L.o_tile_cnt);
    ws.tile_off = (int *)(p + L.o_tile_off);
    ws.fg_bits = (uint32_t *)(p + L.o_bits);
    ws.pix_xy = (uint32_t *)(p + L.o_xy);
    ws.pix_dir = (float2 *)(p + L.o_dir);
    ws.hyp = (float2 *)(p + L.o_hyp);
    ws.pix_pq = (float4 *)(p + L.o_pq);
    ws.fin_ticket = (unsigned *)(p + L.o_ticket);
    ws.fin_part = (double *)(p + L.o_part);
    ws.img_rank = (int *)(p + L.o_rank);
    ws.img_org = (uint32_t *)(p + L.o_org);
    ws.tcq = (uint4 *)(p + L.o_tcq);
    ws.tcq_cnt = (int *)(p + L.o_tcq_cnt);
    ws.tcq_slices = PVV_TCQ_SLICES;
    ws.tcq_slice_cap = L.tcq_slice_cap;
    ws.want_rank = 0;
    ws.idxs_used = nullptr;
    ws.cap = L.cap;
    ws.ntiles = L.ntiles;
    ws.hn_pad = L.hn_pad;
    return ws;
}

pvv_dims make_dims(const pvv_job *j)
{
    pvv_dims d;
    d.b = j->b; d.h = j->h; d.w = j->w; d.vn = j->vn; d.hn = j->hn_total;
    d.fg_rule = j->fg_rule; d.mask_dtype = j->mask_dtype;
    d.min_num = j->min_num; d.max_num = j->max_num;
    d.thresh = j->inlier_thresh;
    d.seed = j->seed; d.image_id_base = j->image_id_base;
    for (int i = 0; i < 3; ++i) d.ms[i] = j->mask_strides[i];
    d.seg_c = j->seg_classes; d.seg_cs = j->seg_class_stride;
    for (int i = 0; i < 5; ++i) d.vs[i] = j->vertex_strides[i];
    return d;
}

// optional per-stage device timing (pvv_debug.stage_ms): events on the job's stream, read back after a sync
struct stage_timer {
    cudaEvent_t ev[5];
    float *out;
    cudaStream_t st;
    int n;
    stage_timer(const pvv_debug *dbg, cudaStream_t s) : out(dbg ? dbg->stage_ms : nullptr), st(s), n(0)
    {
        if (out) for (int i = 0; i < 5; ++i) cudaEventCreate(&ev[i]);
    }
    void mark() { if (out && n < 5) cudaEventRecord(ev[n++], st); }
    void finish()
    {
        if (!out) return;
        cudaEventSynchronize(ev[n - 1]);
        for (int i = 0; i + 1 < n; ++i) cudaEventElapsedTime(&out[i], ev[i], ev[i + 1]);
        cudaEventElapsedTime(&out[4], ev[0], ev[n - 1]);
        for (int i = 0; i < 5; ++i) cudaEventDestroy(ev[i]);
        out = nullptr;
    }
    ~stage_timer() { if (out) for (int i = 0; i < 5; ++i) cudaEventDestroy(ev[i]); }
};

// shared front half of pvv_v3 / pvv_cov: compaction -> hypotheses -> fused counting
int run_front(const pvv_job *j, const pvv_debug *dbg, pvv_dims *d_out, pvv_ws *ws_out, int *launches, stage_timer &tm)
{
    const int v = validate(j);
    if (v != PVV_OK) return v;
    const layout L = make_layout(j);
    if (j->workspace_bytes < L.total_bytes || ((uintptr_t)j->workspace & 255)) return PVV_ERR_WORKSPACE;
    if (dbg && (dbg->pix_xy || dbg->pix_dir || dbg->win_inlier) && dbg->pool_capacity != L.cap) return PVV_ERR_INVALID;
    cudaStream_t st = (cudaStream_t)j->stream;
    pvv_ws ws = carve(L, j->workspace);
    if (dbg) ws.idxs_used = dbg->idxs;
    ws.want_rank = dbg && dbg->k3_variant == 11 ? 1 : 0;   // the experimental k_count_tc2 (variant 10) is the only reader
    const pvv_dims d = make_dims(j);
    cudaError_t e = cudaMemsetAsync(j->workspace, 0, L.zero_bytes, st);
    if (e) return fail_cuda(e);
    tm.mark();
    if ((e = pvv_launch_compact(d, j->mask, j->vertex, j->selection, ws, st, launches))) return fail_cuda(e);
    tm.mark();
    const int force_exact = dbg ? dbg->force_exact : 0, k3_variant = dbg ? dbg->k3_variant : 0;
    if ((e = pvv_launch_hypotheses(d, j->idxs, ws, dbg ? dbg->hyp : nullptr, force_exact, k3_variant,
                                   dbg ? dbg->hyp_inject : nullptr, st, launches))) return fail_cuda(e);
    tm.mark();
    if ((e = pvv_launch_count(d, ws, force_exact, k3_variant, st, launches))) return fail_cuda(e);
    tm.mark();
    *d_out = d;
    *ws_out = ws;
    return PVV_OK;
}

}  // namespace

extern "C" {

const char *pvv_version(void) { return "pvvote 0.1 (sm_100a)"; }
int pvv_last_launch_count(void) { return t_last_launches; }
int pvv_last_cuda_error(void) { return t_last_cuda_error; }

int64_t pvv_default_pool_capacity(const pvv_job *job)
{
    if (!job || job->b <= 0 || job->h <= 0 || job->w <= 0) return 0;
    return pool_alloc(job);
}

size_t pvv_workspace_bytes(const pvv_job *job)
{
    if (!job || job->b <= 0 || job->h <= 0 || job->w <= 0 || job->vn <= 0 || job->hn_total <= 0) return 0;
    return make_layout(job).total_bytes;
}

static int run_v3(const pvv_job *job, float *out_kp, float *out_conf, const pvv_debug *dbg)
{
    if (!out_kp || !job) return PVV_ERR_INVALID;
    int launches = 0;
    pvv_dims d;
    pvv_ws ws;
    stage_timer tm(dbg, (cudaStream_t)job->stream);
    const int r = run_front(job, dbg, &d, &ws, &launches, tm);
    if (r != PVV_OK) return r;
    cudaStream_t st = (cudaStream_t)job->stream;
    cudaError_t e;
    if ((e = pvv_launch_v3_finalize(d, ws, out_kp, out_conf, dbg, st, &launches))) return fail_cuda(e);
    tm.mark();
    tm.finish();
    if ((e = export_status(job, ws))) return fail_cuda(e);
    if ((e = pvv_launch_debug_export(d, ws, dbg, st, &launches))) return fail_cuda(e);
    t_last_launches = launches;
    return PVV_OK;
}

int pvv_v3(const pvv_job *job, float confidence, int max_iter, float *out_kp, const pvv_debug *dbg)
{
    (void)confidence; (void)max_iter;  // unobservable in the reference: every round re-votes the same idxs (:553-582)
    return run_v3(job, out_kp, nullptr, dbg);
}

size_t pvv_static_workspace_bytes(void) { return (size_t)PVV_STATIC_WS_BYTES; }

int pvv_v3_noalloc(const pvv_job *job, float confidence, int max_iter, float *out_kp)
{
    (void)confidence; (void)max_iter;
    if (!job || !out_kp) return PVV_ERR_INVALID;
    // device address of the static scratch: looked up once per device, with that device current (the caller's)
    constexpr int kMaxDev = 64;
    static void *s_ptr[kMaxDev];   // written once per device with the same value by whoever gets there first: benign
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail_cuda(e);
    if (dev < 0 || dev >= kMaxDev) return PVV_ERR_UNSUPPORTED;
    void *ptr = __atomic_load_n(&s_ptr[dev], __ATOMIC_ACQUIRE);
    if (!ptr) {
        if ((e = cudaGetSymbolAddress(&ptr, g_pvv_static_ws)) != cudaSuccess || !ptr) return fail_cuda(e);
        __atomic_store_n(&s_ptr[dev], ptr, __ATOMIC_RELEASE);
    }
    pvv_job j = *job;
    j.workspace = ptr;
    j.workspace_bytes = (size_t)PVV_STATIC_WS_BYTES;
    if (j.b <= 0 || j.h <= 0 || j.w <= 0 || j.vn <= 0 || j.hn_total <= 0) return PVV_ERR_INVALID;
    if (pvv_workspace_bytes(&j) > (size_t)PVV_STATIC_WS_BYTES) return PVV_ERR_WORKSPACE;   // beyond the static maximum
    return run_v3(&j, out_kp, nullptr, nullptr);
}

int pvv_v5(const pvv_job *job, float confidence, int max_iter, float *out_kp, float *out_conf, const pvv_debug *dbg)
{
    (void)confidence; (void)max_iter;  // as in v3: idxs drawn once at :805, identical rounds
    if (!out_conf) return PVV_ERR_INVALID;
    return run_v3(job, out_kp, out_conf, dbg);
}

int pvv_hypotheses(const pvv_job *job, float *out_hyp, int32_t *out_counts, const pvv_debug *dbg)
{
    if (!out_hyp || !out_counts || !job) return PVV_ERR_INVALID;
    int launches = 0;
    pvv_dims d;
    pvv_ws ws;
    stage_timer tm(nullptr, (cudaStream_t)job->stream);
    pvv_debug local;
    memset(&local, 0, sizeof(local));
    if (dbg) local = *dbg;
    local.stage_ms = nullptr;
    local.hyp = out_hyp;   // K2 writes the reference layout [b,hn,vn,2] directly
    const int r = run_front(job, &local, &d, &ws, &launches, tm);
    if (r != PVV_OK) return r;
    cudaStream_t st = (cudaStream_t)job->stream;
    cudaError_t e;
    if ((e = pvv_launch_export_counts(d, ws, out_counts, st, &launches))) return fail_cuda(e);
    if ((e = export_status(job, ws))) return fail_cuda(e);
    local.counts = nullptr;
    if (dbg && (e = pvv_launch_debug_export(d, ws, &local, st, &launches))) return fail_cuda(e);
    t_last_launches = launches;
    return PVV_OK;
}

int pvv_motion(const pvv_job *job, float *out_kp)
{
    if (!out_kp || !job) return PVV_ERR_INVALID;
    pvv_job j = *job;
    j.fg_rule = PVV_FG_BYTE;
    j.min_num = 1;                 // an image without foreground is "skipped": zeros
    j.max_num = 0x7fffffff;        // no subsampling
    j.idxs = nullptr; j.selection = nullptr;
    const int v = validate(&j);
    if (v != PVV_OK) return v;
    const layout L = make_layout(&j);
    if (j.workspace_bytes < L.total_bytes || ((uintptr_t)j.workspace & 255)) return PVV_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)j.stream;
    pvv_ws ws = carve(L, j.workspace);
    const pvv_dims d = make_dims(&j);
    int launches = 0;
    cudaError_t e = cudaMemsetAsync(j.workspace, 0, L.zero_bytes, st);
    if (e) return fail_cuda(e);
    if ((e = pvv_launch_compact(d, j.mask, j.vertex, nullptr, ws, st, &launches))) return fail_cuda(e);
    if ((e = pvv_launch_motion_mean(d, ws, out_kp, st, &launches))) return fail_cuda(e);
    if ((e = export_status(&j, ws))) return fail_cuda(e);
    t_last_launches = launches;
    return PVV_OK;
}

int pvv_cov(const pvv_job *job, const float *mean, float *out_cov, const pvv_debug *dbg)
{
    if (!mean || !out_cov) return PVV_ERR_INVALID;
    int launches = 0;
    pvv_dims d;
    pvv_ws ws;
    if (!job) return PVV_ERR_INVALID;
    stage_timer tm(dbg, (cudaStream_t)job->stream);
    const int r = run_front(job, dbg, &d, &ws, &launches, tm);
    if (r != PVV_OK) return r;
    cudaStream_t st = (cudaStream_t)job->stream;
    cudaError_t e;
    if ((e = pvv_launch_cov_finalize(d, ws, mean, out_cov, st, &launches))) return fail_cuda(e);
    tm.mark();
    tm.finish();
    if ((e = export_status(job, ws))) return fail_cuda(e);
    if ((e = pvv_launch_debug_export(d, ws, dbg, st, &launches))) return fail_cuda(e);
    t_last_launches = launches;
    return PVV_OK;
}

}  // extern "C"
