// pvv_host.cu -- host-buffer entry points (pvv_v3_host / pvv_cov_host), the library's draw streams materialised
// (pvv_make_idxs / pvv_make_selection) and the FP32 roofline probe.  Built only on the public C ABI.
//
// Host path: the mask is copied to the device; a PINNED vertex tensor is not copied at all -- the scatter kernel
// of K1 reads the foreground vectors straight out of host memory over PCIe (UVA), so only ~fg/(h*w) of the
// [b,h,w,vn,2] tensor crosses the bus.  A pageable vertex tensor is staged with one full copy.
#include <stdlib.h>
#include <string.h>

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

struct host_plan {
    size_t dev_ws, o_mask, o_vertex, o_idxs, o_sel, o_mean, o_out, o_tn, o_status, total;
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
            float *out_dev, int *tn_dev, int32_t *status_dev)
{
    pvv_job sj = dj;
    sj.b = i1 - i0;
    sj.status = status_dev;
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
    const int nsub = (mask_pinned && P.vertex_zero_copy && separable && j->b >= 8) ? 4 : 1;

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

    if (nsub == 1) {
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
