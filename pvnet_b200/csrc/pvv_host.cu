// pvv_host.cu -- host-buffer entry points (pvv_v3_host / pvv_cov_host), the library's draw streams materialised
// (pvv_make_idxs / pvv_make_selection) and the FP32 roofline probe.  Built only on the public C ABI.
//
// Host path: the mask is copied to the device; a PINNED vertex tensor is not copied at all -- the scatter kernel
// of K1 reads the foreground vectors straight out of host memory over PCIe (UVA), so only ~fg/(h*w) of the
// [b,h,w,vn,2] tensor crosses the bus.  A pageable vertex tensor is staged with one full copy.
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <thread>
#include <vector>

#include "pvv_internal.cuh"

namespace {

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

size_t mask_elem(int dt) { return dt == PVV_MASK_U8 ? 1 : dt == PVV_MASK_I64 ? 8 : 4; }

// number of elements spanned by a strided view starting at its first element (strides must be >= 0)
long long span_elems(const int *sizes, const int64_t *strides, int n)
{
    long long s = 1;
    for (int i = 0; i < n; ++i) {
        if (strides[i] < 0) return -1;
        s += (long long)(sizes[i] - 1) * strides[i];
    }
    return s;
}

bool is_pinned(const void *p, void **dev_ptr)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    if (a.type == cudaMemoryTypeHost && a.devicePointer) { *dev_ptr = a.devicePointer; return true; }
    return false;
}

// ---- host-side narrowing of wide masks --------------------------------------------------------------------------
// torch.argmax hands the layer an int64 mask (tools/train_linemod.py:254): 8 bytes per pixel of which 7 are zero, 79 % of
// the bytes a host-buffer call of configs[1] used to push over PCIe.  A few host threads narrow it to one byte per pixel
// (the value the foreground rule looks at: the low byte for `.byte()`, 0/1 for `== 1`) into a pinned staging buffer,
// sub-batch by sub-batch, while the previous sub-batch is already crossing the bus and being voted on.
template <typename T>
void narrow_range(const T *src, uint8_t *dst, size_t n, int fg_rule)
{
    if (fg_rule == PVV_FG_EQ1) { for (size_t i = 0; i < n; ++i) dst[i] = src[i] == (T)1 ? 1 : 0; }
    else { for (size_t i = 0; i < n; ++i) dst[i] = (uint8_t)(long long)src[i]; }   // float: truncation toward zero, like .byte()
}

// int64 (what torch.argmax produces): eight pixels per step, the low bytes (or the == 1 flags) gathered into one 64-bit
// store -- ~3 scalar operations per pixel, so a core narrows at memory speed without any ISA-specific code
template <>
void narrow_range<long long>(const long long *src, uint8_t *dst, size_t n, int fg_rule)
{
    const uint64_t *w = (const uint64_t *)src;
    size_t i = 0;
    if (fg_rule == PVV_FG_EQ1) {
        for (; i + 8 <= n; i += 8) {
            uint64_t o = 0;
            for (int k = 0; k < 8; ++k) o |= (uint64_t)(w[i + k] == 1u) << (8 * k);
            memcpy(dst + i, &o, 8);
        }
        for (; i < n; ++i) dst[i] = w[i] == 1u ? 1 : 0;
    } else {
        for (; i + 8 <= n; i += 8) {
            uint64_t o = 0;
            for (int k = 0; k < 8; ++k) o |= (w[i + k] & 0xffu) << (8 * k);
            memcpy(dst + i, &o, 8);
        }
        for (; i < n; ++i) dst[i] = (uint8_t)w[i];
    }
}

void narrow_dispatch(const void *src, int dtype, uint8_t *dst, size_t i0, size_t n, int fg_rule)
{
    switch (dtype) {
        case PVV_MASK_I64: narrow_range((const long long *)src + i0, dst + i0, n, fg_rule); break;
        case PVV_MASK_I32: narrow_range((const int32_t *)src + i0, dst + i0, n, fg_rule); break;
        default: narrow_range((const float *)src + i0, dst + i0, n, fg_rule); break;
    }
}

// pinned staging buffer, one per host thread that calls the host entry points (grown on demand, kept for the thread's
// life: cudaHostAlloc costs milliseconds, a call costs one)
struct pinned_stage {
    void *p = nullptr;
    size_t n = 0;
    ~pinned_stage() { if (p) cudaFreeHost(p); }
    void *get(size_t bytes)
    {
        if (bytes > n) {
            if (p) cudaFreeHost(p);
            p = nullptr; n = 0;
            if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); p = nullptr; return nullptr; }
            n = bytes;
        }
        return p;
    }
};
thread_local pinned_stage t_stage;

std::atomic<int> g_host_workers(-1);   // pvv_set_host_workers: < 0 auto, 0 narrowing off, > 0 that many

struct host_plan {
    size_t dev_ws, sub_ws, o_mask, o_vertex, o_idxs, o_sel, o_mean, o_out, o_tn, o_status, total;
    long long mask_span, vertex_span;
    bool vertex_zero_copy;
    void *vertex_dev;
};

int plan_host(const pvv_job *j, host_plan *P)
{
    if (!j || !j->mask || !j->vertex || j->b <= 0 || j->h <= 0 || j->w <= 0 || j->vn <= 0 || j->hn_total <= 0)
        return PVV_ERR_INVALID;
    const int ms[3] = {j->b, j->h, j->w};
    const int vs[5] = {j->b, j->h, j->w, j->vn, 2};
    P->mask_span = span_elems(ms, j->mask_strides, 3);
    P->vertex_span = span_elems(vs, j->vertex_strides, 5);
    if (P->mask_span < 0 || P->vertex_span < 0) return PVV_ERR_INVALID;
    P->vertex_dev = nullptr;
    P->vertex_zero_copy = is_pinned(j->vertex, &P->vertex_dev);
    size_t o = align_up(pvv_workspace_bytes(j), 256);
    {   // the narrowing path votes on up to 4 sub-batches on two streams at once: each needs a workspace of its own
        const int nsub = j->b >= 8 ? 4 : j->b >= 2 ? 2 : 1;
        pvv_job sj = *j;
        sj.b = (j->b + nsub - 1) / nsub;
        sj.pool_capacity = 0;                       // sub-batches of a narrowed call use the default pool bound
        P->sub_ws = align_up(pvv_workspace_bytes(&sj), 256);
        if ((size_t)nsub * P->sub_ws > o) o = (size_t)nsub * P->sub_ws;
    }
    P->dev_ws = o;
    auto take = [&](size_t bytes) { const size_t at = o; o = align_up(o + bytes, 256); return at; };
    P->o_mask = take((size_t)P->mask_span * mask_elem(j->mask_dtype));
    P->o_vertex = take(P->vertex_zero_copy ? 0 : (size_t)P->vertex_span * sizeof(float));
    P->o_idxs = take(j->idxs ? sizeof(int32_t) * (size_t)j->b * j->hn_total * j->vn * 2 : 0);
    P->o_sel = take(j->selection ? sizeof(float) * (size_t)j->b * j->h * j->w : 0);
    P->o_mean = take(sizeof(float) * (size_t)j->b * j->vn * 2);
    P->o_out = take(sizeof(float) * (size_t)j->b * j->vn * 4);
    P->o_tn = take(sizeof(int) * (size_t)j->b);
    P->o_status = take(sizeof(int32_t) * 4);   // one PVV_FLAG_* word per sub-batch
    P->total = o;
    return PVV_OK;
}

// One sub-batch [i0, i1) of the job on the compute stream; device-side inputs already point at the full arrays.
int run_sub(const pvv_job &dj, int i0, int i1, bool cov, float confidence, int max_iter, const float *mean_dev,
            float *out_dev, int *tn_dev, int32_t *status_dev, cudaStream_t stream = nullptr, void *sub_ws = nullptr,
            size_t sub_ws_bytes = 0)
{
    pvv_job sj = dj;
    sj.b = i1 - i0;
    sj.status = status_dev;
    if (sub_ws) { sj.stream = stream; sj.workspace = sub_ws; sj.workspace_bytes = sub_ws_bytes; }
    sj.mask = (const char *)dj.mask + (size_t)i0 * dj.mask_strides[0] * mask_elem(dj.mask_dtype);
    sj.vertex = dj.vertex + (long long)i0 * dj.vertex_strides[0];
    if (dj.idxs) sj.idxs = dj.idxs + (size_t)i0 * dj.hn_total * dj.vn * 2;
    if (dj.selection) sj.selection = dj.selection + (size_t)i0 * dj.h * dj.w;
    sj.image_id_base = dj.image_id_base + i0;
    pvv_debug dbg;
    memset(&dbg, 0, sizeof(dbg));
    dbg.tn = tn_dev + i0;
    if (cov) return pvv_cov(&sj, mean_dev + (size_t)i0 * dj.vn * 2, out_dev + (size_t)i0 * dj.vn * 4, &dbg);
    return pvv_v3(&sj, confidence, max_iter, out_dev + (size_t)i0 * dj.vn * 2, &dbg);
}

// Host-buffer call.  With pinned inputs the batch is cut into up to 4 sub-batches: the mask of sub-batch s+1 crosses
// PCIe on an internal copy stream while sub-batch s is compacted (its foreground vectors gathered in place over PCIe)
// and voted on the caller's stream, so the bus and the SMs work at the same time.
int run_host(const pvv_job *j, bool cov, float confidence, int max_iter, const float *mean_host, float *out_host,
             int64_t *h2d_bytes, int64_t *d2h_bytes)
{
    host_plan P;
    int r = plan_host(j, &P);
    if (r != PVV_OK) return r;
    if (!j->workspace || j->workspace_bytes < P.total || !out_host) return PVV_ERR_WORKSPACE;
    if (cov && !mean_host) return PVV_ERR_INVALID;
    if (j->mask_dtype == PVV_MASK_LOGITS_F32) return PVV_ERR_UNSUPPORTED;  // logits live on the device, next to the network
    cudaStream_t st = (cudaStream_t)j->stream;
    char *base = (char *)j->workspace;
    int64_t h2d = 0, d2h = 0;
    cudaError_t e;
    pvv_job dj = *j;
    dj.workspace_bytes = P.dev_ws;
    dj.mask = base + P.o_mask;
    const size_t esz = mask_elem(j->mask_dtype);
    const size_t mbytes = (size_t)P.mask_span * esz;
    h2d += (int64_t)mbytes;
    void *pin_dummy = nullptr;
    const bool mask_pinned = is_pinned(j->mask, &pin_dummy);
    // the mask is copied per sub-batch only when images are laid out one after the other (stride[0] covers an image)
    const bool separable = j->mask_strides[0] >= (int64_t)(j->h - 1) * j->mask_strides[1] + (int64_t)(j->w - 1) * j->mask_strides[2] + 1;
    int nsub = (mask_pinned && P.vertex_zero_copy && separable && j->b >= 8) ? 4 : 1;
    // wide, contiguous masks are narrowed to one byte per pixel by host threads (pinned or pageable alike)
    const size_t npx = (size_t)j->h * j->w;
    const bool contiguous = j->mask_strides[2] == 1 && j->mask_strides[1] == j->w && j->mask_strides[0] == (int64_t)npx;
    uint8_t *stage = nullptr;
    const int workers_cfg = g_host_workers.load(std::memory_order_relaxed);
    if (workers_cfg != 0 && esz > 1 && contiguous && P.vertex_zero_copy && npx * j->b >= (size_t)1 << 18)
        stage = (uint8_t *)t_stage.get(npx * j->b);
    const bool narrow = stage != nullptr;
    if (narrow) {
        nsub = j->b >= 8 ? 4 : j->b >= 2 ? 2 : 1;
        dj.mask_dtype = PVV_MASK_U8;           // the staged bytes already are what the foreground rule looks at
        dj.mask_strides[0] = (int64_t)npx; dj.mask_strides[1] = j->w; dj.mask_strides[2] = 1;
        h2d += (int64_t)(npx * j->b) - (int64_t)mbytes;
    }

    if (P.vertex_zero_copy) {
        dj.vertex = (const float *)P.vertex_dev;
    } else {
        const size_t vbytes = (size_t)P.vertex_span * sizeof(float);
        if ((e = cudaMemcpyAsync(base + P.o_vertex, j->vertex, vbytes, cudaMemcpyHostToDevice, st))) return PVV_ERR_CUDA;
        dj.vertex = (const float *)(base + P.o_vertex);
        h2d += (int64_t)vbytes;
    }
    if (j->idxs) {
        const size_t n = sizeof(int32_t) * (size_t)j->b * j->hn_total * j->vn * 2;
        if ((e = cudaMemcpyAsync(base + P.o_idxs, j->idxs, n, cudaMemcpyHostToDevice, st))) return PVV_ERR_CUDA;
        dj.idxs = (const int32_t *)(base + P.o_idxs);
        h2d += (int64_t)n;
    }
    if (j->selection) {
        const size_t n = sizeof(float) * (size_t)j->b * j->h * j->w;
        if ((e = cudaMemcpyAsync(base + P.o_sel, j->selection, n, cudaMemcpyHostToDevice, st))) return PVV_ERR_CUDA;
        dj.selection = (const float *)(base + P.o_sel);
        h2d += (int64_t)n;
    }
    const float *mean_dev = (const float *)(base + P.o_mean);
    if (cov) {
        const size_t n = sizeof(float) * (size_t)j->b * j->vn * 2;
        if ((e = cudaMemcpyAsync(base + P.o_mean, mean_host, n, cudaMemcpyHostToDevice, st))) return PVV_ERR_CUDA;
        h2d += (int64_t)n;
    }
    float *out_dev = (float *)(base + P.o_out);
    int *tn_dev = (int *)(base + P.o_tn);
    int32_t *status_dev = (int32_t *)(base + P.o_status);
    if ((e = cudaMemsetAsync(status_dev, 0, sizeof(int32_t) * 4, st))) return PVV_ERR_CUDA;
    const size_t out_bytes = sizeof(float) * (size_t)j->b * j->vn * (cov ? 4 : 2);

    if (narrow) {
        // workers narrow 64 K-pixel pieces in order; the calling thread hands every finished sub-batch to the copy stream
        // and enqueues its voting on the caller's stream, so host narrowing, PCIe and the SMs overlap
        const size_t piece = (size_t)1 << 16;
        size_t sub_lo[5];
        for (int s = 0; s <= nsub; ++s) sub_lo[s] = (size_t)((long long)j->b * s / nsub) * npx;
        size_t piece_base[5];
        piece_base[0] = 0;
        for (int s = 0; s < nsub; ++s) piece_base[s + 1] = piece_base[s] + (sub_lo[s + 1] - sub_lo[s] + piece - 1) / piece;
        std::atomic<size_t> next(0);
        std::atomic<size_t> done[4];
        for (int s = 0; s < 4; ++s) done[s].store(0);
        const int fg_rule = j->fg_rule, dtype = j->mask_dtype;
        const void *src = j->mask;
        auto work = [&] {
            for (;;) {
                const size_t c = next.fetch_add(1);
                if (c >= piece_base[nsub]) return;
                int s = 0;
                while (c >= piece_base[s + 1]) ++s;
                const size_t i0 = sub_lo[s] + (c - piece_base[s]) * piece;
                const size_t n = i0 + piece <= sub_lo[s + 1] ? piece : sub_lo[s + 1] - i0;
                narrow_dispatch(src, dtype, stage, i0, n, fg_rule);
                done[s].fetch_add(1, std::memory_order_release);
            }
        };
        unsigned hw = std::thread::hardware_concurrency();
#ifndef PVV_HOST_THREADS_DIV
#define PVV_HOST_THREADS_DIV 2   // measured on the 16-thread host of a B200 box: 8 workers beat 12 and 4 (profiles/r02_e2e_host_threads.log)
#endif
        int nthreads = workers_cfg > 0 ? workers_cfg : (int)(hw / PVV_HOST_THREADS_DIV);
        nthreads = nthreads < 1 ? 1 : nthreads > 16 ? 16 : nthreads;
        if ((size_t)nthreads > piece_base[nsub]) nthreads = (int)piece_base[nsub];
        std::vector<std::thread> pool;
        for (int t = 0; t < nthreads; ++t) pool.emplace_back(work);
        // one copy stream and two voting streams: the compaction of sub-batch s + 1 gathers its foreground vectors over PCIe
        // (bus-bound) while sub-batch s is being voted on (SM-bound); every sub-batch has its own workspace
        cudaStream_t cs = nullptr, ks[2] = {nullptr, nullptr};
        cudaEvent_t ev[7] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
        bool ok = cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking) == cudaSuccess &&
                  cudaStreamCreateWithFlags(&ks[0], cudaStreamNonBlocking) == cudaSuccess &&
                  cudaStreamCreateWithFlags(&ks[1], cudaStreamNonBlocking) == cudaSuccess;
        for (int i = 0; ok && i < 7; ++i) ok = cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming) == cudaSuccess;
        // everything the caller's stream did before (input copies above, earlier use of the workspace) comes first
        ok = ok && cudaEventRecord(ev[4], st) == cudaSuccess && cudaStreamWaitEvent(cs, ev[4], 0) == cudaSuccess &&
             cudaStreamWaitEvent(ks[0], ev[4], 0) == cudaSuccess && cudaStreamWaitEvent(ks[1], ev[4], 0) == cudaSuccess;
        r = PVV_OK;
        for (int s = 0; s < nsub; ++s) {
            const size_t want = piece_base[s + 1] - piece_base[s];
            while (done[s].load(std::memory_order_acquire) < want) std::this_thread::yield();
            if (!ok || r != PVV_OK) continue;   // keep draining the workers; nothing more is enqueued
            const int i0 = (int)((long long)j->b * s / nsub), i1 = (int)((long long)j->b * (s + 1) / nsub);
            cudaStream_t k = ks[s & 1];
            ok = cudaMemcpyAsync(base + P.o_mask + sub_lo[s], stage + sub_lo[s], sub_lo[s + 1] - sub_lo[s], cudaMemcpyHostToDevice, cs) ==
                     cudaSuccess &&
                 cudaEventRecord(ev[s], cs) == cudaSuccess && cudaStreamWaitEvent(k, ev[s], 0) == cudaSuccess;
            if (ok) {
                pvv_job sub = dj;
                sub.pool_capacity = 0;
                r = run_sub(sub, i0, i1, cov, confidence, max_iter, mean_dev, out_dev, tn_dev, status_dev + s, k,
                            base + (size_t)s * P.sub_ws, P.sub_ws);
            }
        }
        for (auto &t : pool) t.join();
        // the caller's stream continues after both voting streams
        for (int i = 0; ok && i < 2; ++i)
            ok = cudaEventRecord(ev[5 + i], ks[i]) == cudaSuccess && cudaStreamWaitEvent(st, ev[5 + i], 0) == cudaSuccess;
        for (int i = 0; i < 7; ++i) if (ev[i]) cudaEventDestroy(ev[i]);
        if (cs) cudaStreamDestroy(cs);
        for (int i = 0; i < 2; ++i) if (ks[i]) cudaStreamDestroy(ks[i]);
        if (r != PVV_OK) return r;
        if (!ok) { cudaGetLastError(); return PVV_ERR_CUDA; }
    } else if (nsub == 1) {
        if ((e = cudaMemcpyAsync(base + P.o_mask, j->mask, mbytes, cudaMemcpyHostToDevice, st))) return PVV_ERR_CUDA;
        r = run_sub(dj, 0, j->b, cov, confidence, max_iter, mean_dev, out_dev, tn_dev, status_dev);
        if (r != PVV_OK) return r;
    } else {
        cudaStream_t cs = nullptr;
        cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
        bool ok = cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking) == cudaSuccess;
        for (int i = 0; ok && i < 5; ++i) ok = cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming) == cudaSuccess;
        // the copy stream starts after whatever the caller's stream did before (the workspace may still be in use)
        ok = ok && cudaEventRecord(ev[4], st) == cudaSuccess && cudaStreamWaitEvent(cs, ev[4], 0) == cudaSuccess;
        r = PVV_OK;
        for (int s = 0; ok && s < nsub; ++s) {
            const int i0 = (int)((long long)j->b * s / nsub), i1 = (int)((long long)j->b * (s + 1) / nsub);
            const size_t o0 = (size_t)i0 * j->mask_strides[0] * esz;
            const size_t o1 = s + 1 == nsub ? mbytes : (size_t)i1 * j->mask_strides[0] * esz;
            ok = cudaMemcpyAsync(base + P.o_mask + o0, (const char *)j->mask + o0, o1 - o0, cudaMemcpyHostToDevice, cs) ==
                     cudaSuccess &&
                 cudaEventRecord(ev[s], cs) == cudaSuccess;
        }
        for (int s = 0; ok && r == PVV_OK && s < nsub; ++s) {
            const int i0 = (int)((long long)j->b * s / nsub), i1 = (int)((long long)j->b * (s + 1) / nsub);
            ok = cudaStreamWaitEvent(st, ev[s], 0) == cudaSuccess;
            if (ok) r = run_sub(dj, i0, i1, cov, confidence, max_iter, mean_dev, out_dev, tn_dev, status_dev + s);
        }
        // destroying an event / a stream with work in flight is legal: the resources go away when that work completes,
        // and the caller's stream already waits on every copy through the events
        for (int i = 0; i < 5; ++i) if (ev[i]) cudaEventDestroy(ev[i]);
        if (cs) cudaStreamDestroy(cs);
        if (r != PVV_OK) return r;
        if (!ok) { cudaGetLastError(); return PVV_ERR_CUDA; }
    }
    if ((e = cudaMemcpyAsync(out_host, out_dev, out_bytes, cudaMemcpyDeviceToHost, st))) return PVV_ERR_CUDA;
    d2h += (int64_t)out_bytes;
    int32_t status_host[4] = {0, 0, 0, 0};
    if ((e = cudaMemcpyAsync(status_host, status_dev, sizeof(status_host), cudaMemcpyDeviceToHost, st))) return PVV_ERR_CUDA;
    d2h += (int64_t)sizeof(status_host);
    if (P.vertex_zero_copy && h2d_bytes) {
        // account the foreground vectors the scatter kernel pulled over the bus: sum(tn) * vn * 8 bytes
        int *tn_host = (int *)malloc(sizeof(int) * j->b);
        if (!tn_host) return PVV_ERR_INVALID;
        e = cudaMemcpyAsync(tn_host, tn_dev, sizeof(int) * j->b, cudaMemcpyDeviceToHost, st);
        if (!e) e = cudaStreamSynchronize(st);
        long long sum = 0;
        for (int i = 0; i < j->b; ++i) sum += tn_host[i];
        free(tn_host);
        if (e) return PVV_ERR_CUDA;
        h2d += sum * j->vn * 8;
        d2h += (int64_t)sizeof(int) * j->b;
    } else {
        if ((e = cudaStreamSynchronize(st))) return PVV_ERR_CUDA;
    }
    if (h2d_bytes) *h2d_bytes = h2d;
    if (d2h_bytes) *d2h_bytes = d2h;
    // the call has synchronised: a pool overflow (user-set pool_capacity too small) is reported, not silent
    if ((status_host[0] | status_host[1] | status_host[2] | status_host[3]) & PVV_FLAG_POOL_OVERFLOW) return PVV_ERR_POOL_OVERFLOW;
    return PVV_OK;
}

// ---- materialised draw streams ---------------------------------------------------------------------
__global__ void k_make_idxs(unsigned long long seed, long long base, int b, int hn, int vn, const int32_t *tn,
                            int32_t *idxs)
{
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (long long)b * hn * vn) return;
    const int vi = (int)(e % vn);
    const long long r = e / vn;
    const int hi = (int)(r % hn), img = (int)(r / hn);
    int t0 = 0, t1 = 0;
    const int n = tn[img];
    if (n > 0) pvv_draw_pair(seed, base + img, hi, vi, vn, n, t0, t1);
    idxs[e * 2] = t0;
    idxs[e * 2 + 1] = t1;
}

__global__ void k_make_selection(unsigned long long seed, long long base, int b, int npx, float *sel)
{
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (long long)b * npx) return;
    sel[e] = pvv_draw_uniform(seed, base + e / npx, (uint32_t)(e % npx));
}

// ---- FP32 probe: 8 independent FFMA chains per thread ----------------------------------------------
__global__ void __launch_bounds__(256) k_ffma_probe(float *out, int iters, float a, float b)
{
    float r0 = threadIdx.x, r1 = r0 + 1, r2 = r0 + 2, r3 = r0 + 3, r4 = r0 + 4, r5 = r0 + 5, r6 = r0 + 6, r7 = r0 + 7;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            r0 = fmaf(r0, a, b); r1 = fmaf(r1, a, b); r2 = fmaf(r2, a, b); r3 = fmaf(r3, a, b);
            r4 = fmaf(r4, a, b); r5 = fmaf(r5, a, b); r6 = fmaf(r6, a, b); r7 = fmaf(r7, a, b);
        }
    }
    const float s = r0 + r1 + r2 + r3 + r4 + r5 + r6 + r7;
    if (s == 12345.678f) out[0] = s;  // never true in practice; keeps the chains alive
}

// packed variant: fma.rn.f32x2 (FFMA2 in SASS), two FMAs per instruction and thread
__global__ void __launch_bounds__(256) k_ffma2_probe(float *out, int iters, float a, float b)
{
    unsigned long long r[8], av, bv;
    asm("mov.b64 %0, {%1, %1};" : "=l"(av) : "f"(a));
    asm("mov.b64 %0, {%1, %1};" : "=l"(bv) : "f"(b));
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const float x = threadIdx.x + i;
        asm("mov.b64 %0, {%1, %1};" : "=l"(r[i]) : "f"(x));
    }
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
#pragma unroll
            for (int k = 0; k < 8; ++k) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(r[k]) : "l"(av), "l"(bv));
        }
    }
    unsigned long long s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s ^= r[i];
    if (s == 0x123456789ull) out[0] = 1.f;
}

}  // namespace

extern "C" {

int pvv_set_host_workers(int n) { return g_host_workers.exchange(n < -1 ? -1 : n); }

size_t pvv_workspace_bytes_host(const pvv_job *job)
{
    host_plan P;
    if (plan_host(job, &P) != PVV_OK) return 0;
    return P.total;
}

int pvv_v3_host(const pvv_job *job, float confidence, int max_iter, float *out_kp_host, int64_t *h2d_bytes,
                int64_t *d2h_bytes)
{
    return run_host(job, false, confidence, max_iter, nullptr, out_kp_host, h2d_bytes, d2h_bytes);
}

int pvv_cov_host(const pvv_job *job, const float *mean_host, float *out_cov_host, int64_t *h2d_bytes,
                 int64_t *d2h_bytes)
{
    return run_host(job, true, 0.f, 0, mean_host, out_cov_host, h2d_bytes, d2h_bytes);
}

int pvv_make_idxs(uint64_t seed, int64_t image_id_base, int b, int hn_total, int vn, const int32_t *tn, int32_t *idxs,
                  void *stream)
{
    if (!tn || !idxs || b <= 0 || hn_total <= 0 || vn <= 0) return PVV_ERR_INVALID;
    const long long total = (long long)b * hn_total * vn;
    k_make_idxs<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(seed, image_id_base, b, hn_total, vn,
                                                                                    tn, idxs);
    return cudaGetLastError() == cudaSuccess ? PVV_OK : PVV_ERR_CUDA;
}

int pvv_make_selection(uint64_t seed, int64_t image_id_base, int b, int h, int w, float *selection, void *stream)
{
    if (!selection || b <= 0 || h <= 0 || w <= 0) return PVV_ERR_INVALID;
    const long long total = (long long)b * h * w;
    k_make_selection<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(seed, image_id_base, b, h * w,
                                                                                        selection);
    return cudaGetLastError() == cudaSuccess ? PVV_OK : PVV_ERR_CUDA;
}

double pvv_probe_tflops(int kind, int iters, void *stream);
double pvv_fp32_peak_tflops(int iters, void *stream) { return pvv_probe_tflops(0, iters, stream); }

double pvv_probe_tflops(int kind, int iters, void *stream)
{
    cudaStream_t st = (cudaStream_t)stream;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (iters <= 0) iters = 4096;
    float *out = nullptr;
    if (cudaMalloc(&out, sizeof(float)) != cudaSuccess) return 0.0;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int blocks = sms * 8;
    const double per_instr = kind == 1 ? 4.0 : 2.0;
    if (kind == 1) k_ffma2_probe<<<blocks, 256, 0, st>>>(out, 64, 1.0000001f, 1e-9f);
    else k_ffma_probe<<<blocks, 256, 0, st>>>(out, 64, 1.0000001f, 1e-9f);  // warm-up
    double best = 0.0;
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0, st);
        if (kind == 1) k_ffma2_probe<<<blocks, 256, 0, st>>>(out, iters, 1.0000001f, 1e-9f);
        else k_ffma_probe<<<blocks, 256, 0, st>>>(out, iters, 1.0000001f, 1e-9f);
        cudaEventRecord(e1, st);
        if (cudaEventSynchronize(e1) != cudaSuccess) { best = 0.0; break; }
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        const double flop = per_instr * 8 * 16 * (double)iters * 256.0 * blocks;
        const double tf = flop / (ms * 1e-3) / 1e12;
        if (tf > best) best = tf;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(out);
    return best;
}

}  // extern "C"
