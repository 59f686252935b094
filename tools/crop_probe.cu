// Probe (GPU box): how fast do the copy engines move bounding-box crops of a pinned NCHW vertex tensor, against one whole copy?
// build: nvcc -O2 -gencode arch=compute_100a,code=sm_100a tools/crop_probe.cu -o tools/_build/crop_probe
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <atomic>
#include <thread>
#include <vector>
#define CK(x) do { cudaError_t e_ = (x); if (e_) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)
// SM pull: warps read window rows from mapped pinned memory (coalesced 4-byte loads, ROWS rows in flight per warp)
struct win { int x0, y0, wx, hy; };
template <int ROWS>
__global__ void k_pull(const float *__restrict__ src, float *__restrict__ dst, const win *wins, int b, int planes, int h, int w)
{
    const int warps = gridDim.x * (blockDim.x >> 5), gw = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    // rows of all images are dealt round-robin in groups of ROWS
    long long total = 0;
    for (int i = 0; i < b; ++i) total += (long long)planes * wins[i].hy;
    for (long long r0 = (long long)gw * ROWS; r0 < total; r0 += (long long)warps * ROWS) {
        float v[ROWS][4];
        long long offs[ROWS];
        int wxs[ROWS];
#pragma unroll
        for (int k = 0; k < ROWS; ++k) {
            long long r = r0 + k;
            offs[k] = -1; wxs[k] = 0;
            if (r < total) {
                int i = 0;
                while (r >= (long long)planes * wins[i].hy) { r -= (long long)planes * wins[i].hy; ++i; }
                const int pl = (int)(r / wins[i].hy), y = wins[i].y0 + (int)(r % wins[i].hy);
                offs[k] = (((long long)i * planes + pl) * h + y) * w + wins[i].x0;
                wxs[k] = wins[i].wx;
            }
        }
#pragma unroll
        for (int k = 0; k < ROWS; ++k)
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int x = q * 32 + lane;
                v[k][q] = (offs[k] >= 0 && x < wxs[k]) ? __ldg(src + offs[k] + x) : 0.f;
            }
#pragma unroll
        for (int k = 0; k < ROWS; ++k)
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int x = q * 32 + lane;
                if (offs[k] >= 0 && x < wxs[k]) dst[offs[k] + x] = v[k][q];
            }
    }
}

// ---- the library's pull kernel (pvv_host.cu), pasted for the probe ----
#include <stdint.h>
template <bool PLANAR, bool VEC4>
__global__ void __launch_bounds__(64, 32) k_pull_spans(const uint8_t *__restrict__ mask, const float *__restrict__ src,
                                                       float *__restrict__ dst, int b, int h, int w, int vn, long long vs0,
                                                       long long vs1, long long vs2, long long vs3, long long vs4)
{
    const int lane = threadIdx.x & 31;
    const int nwarps = (int)gridDim.x * 2;
    for (int task = (int)blockIdx.x * 2 + (threadIdx.x >> 5); task < b * h; task += nwarps) {
        const int img = task / h, y = task - img * h;
        const uint8_t *row = mask + (size_t)task * w;
        int first = 0x7fffffff, last = -1;
        if ((w & 3) == 0) {   // rows start on 4-byte boundaries (the mask buffer is 256-byte aligned)
            for (int x = lane * 4; x < w; x += 128) {
                const uint32_t v = *reinterpret_cast<const uint32_t *>(row + x);
                if (v) {
                    first = min(first, x + ((__ffs((int)v) - 1) >> 3));
                    last = max(last, x + 3 - (__clz((int)v) >> 3));
                }
            }
        } else {
            for (int x = lane; x < w; x += 32)
                if (row[x]) { first = min(first, x); last = max(last, x); }
        }
        first = __reduce_min_sync(0xffffffffu, first);
        last = __reduce_max_sync(0xffffffffu, last);
        if (last < 0) continue;
        const long long base = (long long)img * vs0 + (long long)y * vs1;
        if (PLANAR) {
            const float *sb = src + base;
            float *db = dst + base;
            const int planes = 2 * vn;
            for (int p0 = 0; p0 < planes; p0 += 4) {
                int po[4];   // plane offsets inside the image: 32 bits (the host checks the image stride)
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int p = min(p0 + k, planes - 1);
                    po[k] = (int)((long long)(p >> 1) * vs3 + (long long)(p & 1) * vs4);
                }
                if (VEC4) {
                    // 16-byte loads (rows, planes and both tensors are 16-byte aligned, w % 4 == 0): a lane moves 4 pixels, a
                    // warp 128 -- one round trip per plane for a typical span, four planes in flight; up to 3 pixels
                    // either side of the span ride along (they stay inside the row)
                    for (int xs = (first & ~3) + 4 * lane; xs <= last; xs += 128) {
                        float4 v[4];
#pragma unroll
                        for (int k = 0; k < 4; ++k) v[k] = __ldg(reinterpret_cast<const float4 *>(sb + po[k] + xs));
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            if (p0 + k < planes) *reinterpret_cast<float4 *>(db + po[k] + xs) = v[k];
                    }
                    continue;
                }
                for (int xs = first & ~31; xs <= last; xs += 64) {
                    const int xa = xs + lane, xb = xa + 32;
                    const bool oka = xa >= first && xa <= last, okb = xb <= last;   // xb >= first always
                    float va[4], vb[4];
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        va[k] = oka ? __ldg(sb + po[k] + xa) : 0.f;
                        vb[k] = okb ? __ldg(sb + po[k] + xb) : 0.f;
                    }
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        if (p0 + k < planes) {
                            if (oka) db[po[k] + xa] = va[k];
                            if (okb) db[po[k] + xb] = vb[k];
                        }
                }
            }
        } else {
            // one run of floats: [first * vs2, (last + 1) * vs2), vs2 = 2 vn
            const long long r0 = base + (long long)first * vs2, r1 = base + (long long)(last + 1) * vs2;
            for (long long o = (r0 & ~31LL) + lane; o < r1; o += 32 * 8) {
                float v[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) { const long long q = o + 32 * k; v[k] = (q >= r0 && q < r1) ? __ldg(src + q) : 0.f; }
#pragma unroll
                for (int k = 0; k < 8; ++k) { const long long q = o + 32 * k; if (q >= r0 && q < r1) dst[q] = v[k]; }
            }
        }
    }
}


int main(int argc, char **argv)
{
    const int b = 32, planes = 18, h = 480, w = 640;
    const int box = argc > 1 ? atoi(argv[1]) : 112;   // box x box pixels: 112^2 = 4.1 % of the image (3 % disc -> 3.8 % box)
    const size_t n = (size_t)b * planes * h * w;
    float *hp, *dp;
    CK(cudaHostAlloc(&hp, n * 4, cudaHostAllocDefault));
    CK(cudaMalloc(&dp, n * 4));
    for (size_t i = 0; i < n; i += 1024) hp[i] = 1.f;
    cudaStream_t st[4];
    for (int i = 0; i < 4; ++i) CK(cudaStreamCreateWithFlags(&st[i], cudaStreamNonBlocking));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (int nstreams = 1; nstreams <= 4; nstreams *= 2) {
        for (int rep = 0; rep < 3; ++rep) {
            CK(cudaDeviceSynchronize());
            CK(cudaEventRecord(e0, st[0]));
            for (int s = 1; s < nstreams; ++s) CK(cudaStreamWaitEvent(st[s], e0, 0));
            for (int i = 0; i < b; ++i) {
                const int x0 = (37 * i) % (w - box), y0 = (53 * i) % (h - box);
                cudaMemcpy3DParms p = {};
                p.srcPtr = make_cudaPitchedPtr(hp + (size_t)i * planes * h * w, (size_t)w * 4, w, h);
                p.dstPtr = make_cudaPitchedPtr(dp + (size_t)i * planes * h * w, (size_t)w * 4, w, h);
                p.srcPos = make_cudaPos((size_t)x0 * 4, y0, 0);
                p.dstPos = p.srcPos;
                p.extent = make_cudaExtent((size_t)box * 4, box, planes);
                p.kind = cudaMemcpyHostToDevice;
                CK(cudaMemcpy3DAsync(&p, st[i % nstreams]));
            }
            cudaEvent_t ej;
            for (int s = 1; s < nstreams; ++s) { CK(cudaEventCreate(&ej)); CK(cudaEventRecord(ej, st[s])); CK(cudaStreamWaitEvent(st[0], ej, 0)); }
            CK(cudaEventRecord(e1, st[0]));
            CK(cudaEventSynchronize(e1));
            float ms;
            CK(cudaEventElapsedTime(&ms, e0, e1));
            const double mb = (double)b * planes * box * box * 4 / 1e6;
            printf("3-D crops %dx%d x %d planes x %d images, %d stream(s): %.3f ms  %.1f MB  %.1f GB/s\n", box, box, planes, b, nstreams, ms, mb, mb / ms);
        }
    }
    // interleaved layout [b][h][w][18]: one 2-D copy per image, rows of box * 72 bytes
    for (int rep = 0; rep < 3; ++rep) {
        CK(cudaDeviceSynchronize());
        CK(cudaEventRecord(e0, st[0]));
        for (int i = 0; i < b; ++i) {
            const int x0 = (37 * i) % (w - box), y0 = (53 * i) % (h - box);
            const size_t off = ((size_t)i * h * w + (size_t)y0 * w + x0) * planes;
            CK(cudaMemcpy2DAsync(dp + off, (size_t)w * planes * 4, hp + off, (size_t)w * planes * 4, (size_t)box * planes * 4, box, cudaMemcpyHostToDevice, st[0]));
        }
        CK(cudaEventRecord(e1, st[0]));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        const double mb = (double)b * planes * box * box * 4 / 1e6;
        printf("2-D crops (interleaved) rows of %d B: %.3f ms  %.1f MB  %.1f GB/s\n", box * planes * 4, ms, mb, mb / ms);
    }
    // per-plane 2-D copies (18 x 32 calls)
    for (int rep = 0; rep < 2; ++rep) {
        CK(cudaDeviceSynchronize());
        CK(cudaEventRecord(e0, st[0]));
        for (int i = 0; i < b; ++i)
            for (int pl = 0; pl < planes; ++pl) {
                const int x0 = (37 * i) % (w - box), y0 = (53 * i) % (h - box);
                const size_t off = (((size_t)i * planes + pl) * h + y0) * w + x0;
                CK(cudaMemcpy2DAsync(dp + off, (size_t)w * 4, hp + off, (size_t)w * 4, (size_t)box * 4, box, cudaMemcpyHostToDevice, st[0]));
            }
        CK(cudaEventRecord(e1, st[0]));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("2-D per-plane copies (%d calls): %.3f ms\n", b * planes, ms);
    }
    // full-row bands: rows y0..y0+box of every plane, whole width (contiguous per plane): 18 x 32 plain copies in one 2-D call per image
    for (int rep = 0; rep < 2; ++rep) {
        CK(cudaDeviceSynchronize());
        CK(cudaEventRecord(e0, st[0]));
        for (int i = 0; i < b; ++i) {
            const int y0 = (53 * i) % (h - box);
            const size_t off = (size_t)i * planes * h * w + (size_t)y0 * w;
            CK(cudaMemcpy2DAsync(dp + off, (size_t)h * w * 4, hp + off, (size_t)h * w * 4, (size_t)box * w * 4, planes, cudaMemcpyHostToDevice, st[0]));
        }
        CK(cudaEventRecord(e1, st[0]));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        const double mb = (double)b * planes * box * w * 4 / 1e6;
        printf("row bands (full width, %d rows): %.3f ms  %.1f MB  %.1f GB/s\n", box, ms, mb, mb / ms);
    }
    // crops on one stream + a contiguous 9.8 MB copy (the byte masks) on another: do the two overlap?
    {
        unsigned char *hm, *dm;
        const size_t mb = (size_t)b * h * w;
        CK(cudaHostAlloc(&hm, mb, cudaHostAllocDefault));
        CK(cudaMalloc(&dm, mb));
        memset(hm, 1, mb);
        for (int mode = 0; mode < 3; ++mode)
            for (int rep = 0; rep < 2; ++rep) {
                CK(cudaDeviceSynchronize());
                CK(cudaEventRecord(e0, st[0]));
                CK(cudaStreamWaitEvent(st[1], e0, 0));
                for (int i = 0; i < b; ++i) {
                    const int x0 = (37 * i) % (w - box), y0 = (53 * i) % (h - box);
                    cudaMemcpy3DParms p = {};
                    p.srcPtr = make_cudaPitchedPtr(hp + (size_t)i * planes * h * w, (size_t)w * 4, w, h);
                    p.dstPtr = make_cudaPitchedPtr(dp + (size_t)i * planes * h * w, (size_t)w * 4, w, h);
                    p.srcPos = make_cudaPos((size_t)x0 * 4, y0, 0);
                    p.dstPos = p.srcPos;
                    p.extent = make_cudaExtent((size_t)box * 4, box, planes);
                    p.kind = cudaMemcpyHostToDevice;
                    if (mode != 1) CK(cudaMemcpy3DAsync(&p, st[0]));
                    if (mode != 0 && i % 4 == 0) CK(cudaMemcpyAsync(dm + (size_t)i * h * w, hm + (size_t)i * h * w, (size_t)4 * h * w, cudaMemcpyHostToDevice, st[1]));
                }
                cudaEvent_t ej;
                CK(cudaEventCreate(&ej)); CK(cudaEventRecord(ej, st[1])); CK(cudaStreamWaitEvent(st[0], ej, 0));
                CK(cudaEventRecord(e1, st[0]));
                CK(cudaEventSynchronize(e1));
                float ms;
                CK(cudaEventElapsedTime(&ms, e0, e1));
                printf("%s: %.3f ms\n", mode == 0 ? "crops alone" : mode == 1 ? "byte masks alone (8 copies of 1.2 MB)" : "crops + byte masks on two streams", ms);
            }
    }
    {   // SM pull of the same windows (box <= 128 wide)
        win hw_[32], *dw;
        for (int i = 0; i < b; ++i) hw_[i] = {(37 * i) % (w - box), (53 * i) % (h - box), box, box};
        CK(cudaMalloc(&dw, sizeof(hw_)));
        CK(cudaMemcpy(dw, hw_, sizeof(hw_), cudaMemcpyHostToDevice));
        float *hmap;
        CK(cudaHostGetDevicePointer(&hmap, hp, 0));
        for (int grid : {148, 296, 592, 1184})
            for (int rep = 0; rep < 2; ++rep) {
                CK(cudaDeviceSynchronize());
                CK(cudaEventRecord(e0, st[0]));
                if (box <= 128) k_pull<4><<<grid, 256, 0, st[0]>>>(hmap, dp, dw, b, planes, h, w);
                CK(cudaEventRecord(e1, st[0]));
                CK(cudaEventSynchronize(e1));
                CK(cudaGetLastError());
                float ms;
                CK(cudaEventElapsedTime(&ms, e0, e1));
                const double mb = (double)b * planes * box * box * 4 / 1e6;
                if (rep) printf("SM pull <4 rows>, grid %4d: %.3f ms  %.1f MB  %.1f GB/s\n", grid, ms, mb, mb / ms);
            }
        // footprint sweep: CTA size x rows in flight, one CTA per SM
        for (int threads : {64, 128, 256, 512})
            for (int rows = 1; rows <= 4; rows *= 2)
                for (int rep = 0; rep < 2; ++rep) {
                    CK(cudaDeviceSynchronize());
                    CK(cudaEventRecord(e0, st[0]));
                    if (rows == 1) k_pull<1><<<148, threads, 0, st[0]>>>(hmap, dp, dw, b, planes, h, w);
                    else if (rows == 2) k_pull<2><<<148, threads, 0, st[0]>>>(hmap, dp, dw, b, planes, h, w);
                    else k_pull<4><<<148, threads, 0, st[0]>>>(hmap, dp, dw, b, planes, h, w);
                    CK(cudaEventRecord(e1, st[0]));
                    CK(cudaEventSynchronize(e1));
                    float ms;
                    CK(cudaEventElapsedTime(&ms, e0, e1));
                    const double mb = (double)b * planes * box * box * 4 / 1e6;
                    if (rep) printf("SM pull: 148 CTAs x %3d threads, %d rows (x4 loads) in flight: %.3f ms  %.1f GB/s\n", threads, rows, ms, mb / ms);
                }
        for (int grid : {296, 1184})
            for (int rep = 0; rep < 2; ++rep) {
                CK(cudaDeviceSynchronize());
                CK(cudaEventRecord(e0, st[0]));
                if (box <= 128) k_pull<8><<<grid, 256, 0, st[0]>>>(hmap, dp, dw, b, planes, h, w);
                CK(cudaEventRecord(e1, st[0]));
                CK(cudaEventSynchronize(e1));
                float ms;
                CK(cudaEventElapsedTime(&ms, e0, e1));
                const double mb = (double)b * planes * box * box * 4 / 1e6;
                if (rep) printf("SM pull <8 rows>, grid %4d: %.3f ms  %.1f MB  %.1f GB/s\n", grid, ms, mb, mb / ms);
            }
    }
    {   // the library's kernel on a byte mask with box-shaped foreground, 8 images per launch like a sub-batch, and all 32
        unsigned char *hm2, *dm2;
        const size_t mb2 = (size_t)b * h * w;
        CK(cudaHostAlloc(&hm2, mb2, cudaHostAllocDefault));
        CK(cudaMalloc(&dm2, mb2));
        memset(hm2, 0, mb2);
        for (int i = 0; i < b; ++i) {
            const int x0 = (37 * i) % (w - box), y0 = (53 * i) % (h - box);
            for (int y = y0; y < y0 + box; ++y) memset(hm2 + ((size_t)i * h + y) * w + x0, 1, box);
        }
        CK(cudaMemcpy(dm2, hm2, mb2, cudaMemcpyHostToDevice));
        float *hmap;
        CK(cudaHostGetDevicePointer(&hmap, hp, 0));
        const long long vs0 = (long long)planes * h * w, vs1 = w, vs2 = 1, vs3 = 2LL * h * w, vs4 = (long long)h * w;
        for (int per : {32, 8, 4})
            for (int vec = 0; vec < 2; ++vec)
                for (int rep = 0; rep < 2; ++rep) {
                    CK(cudaDeviceSynchronize());
                    CK(cudaEventRecord(e0, st[0]));
                    for (int i0 = 0; i0 < b; i0 += per) {
                        if (vec) k_pull_spans<true, true><<<148, 64, 0, st[0]>>>(dm2 + (size_t)i0 * h * w, hmap + i0 * vs0, dp + i0 * vs0, per, h, w, planes / 2, vs0, vs1, vs2, vs3, vs4);
                        else k_pull_spans<true, false><<<148, 64, 0, st[0]>>>(dm2 + (size_t)i0 * h * w, hmap + i0 * vs0, dp + i0 * vs0, per, h, w, planes / 2, vs0, vs1, vs2, vs3, vs4);
                    }
                    CK(cudaEventRecord(e1, st[0]));
                    CK(cudaEventSynchronize(e1));
                    CK(cudaGetLastError());
                    float ms;
                    CK(cudaEventElapsedTime(&ms, e0, e1));
                    const double mb = (double)b * planes * box * box * 4 / 1e6;
                    if (rep) printf("library pull kernel, %2d images per launch, %s: %.3f ms  %.1f GB/s\n", per, vec ? "16-byte loads" : "4-byte loads", ms, mb / ms);
                }
    }
    {   // the library's pull kernel under load: (a) byte-mask copies by the copy engine at the same time, (b) 8 host threads
        // streaming through host memory at the same time (what the narrowing workers do)
        unsigned char *hm3, *dm3;
        const size_t mb3 = (size_t)b * h * w;
        CK(cudaHostAlloc(&hm3, mb3, cudaHostAllocDefault));
        CK(cudaMalloc(&dm3, mb3));
        memset(hm3, 0, mb3);
        for (int i = 0; i < b; ++i) {
            const int x0 = (37 * i) % (w - box), y0 = (53 * i) % (h - box);
            for (int y = y0; y < y0 + box; ++y) memset(hm3 + ((size_t)i * h + y) * w + x0, 1, box);
        }
        CK(cudaMemcpy(dm3, hm3, mb3, cudaMemcpyHostToDevice));
        float *hmap;
        CK(cudaHostGetDevicePointer(&hmap, hp, 0));
        const long long vs0 = (long long)planes * h * w, vs1 = w, vs2 = 1, vs3 = 2LL * h * w, vs4 = (long long)h * w;
        const size_t big = (size_t)512 << 20;
        unsigned long long *hbig = (unsigned long long *)malloc(big);
        memset(hbig, 1, big);
        for (int mode = 0; mode < 4; ++mode) {
            std::atomic<bool> stop(false);
            std::atomic<unsigned long long> sink(0);
            std::vector<std::thread> pool;
            if (mode & 2)
                for (int t = 0; t < 8; ++t)
                    pool.emplace_back([&, t] {
                        unsigned long long acc = 0;
                        const size_t n = big / 8 / 8;
                        while (!stop.load(std::memory_order_relaxed))
                            for (size_t i = 0; i < n; ++i) acc += hbig[t * n + i];
                        sink += acc;
                    });
            for (int rep = 0; rep < 3; ++rep) {
                CK(cudaDeviceSynchronize());
                CK(cudaEventRecord(e0, st[0]));
                CK(cudaStreamWaitEvent(st[1], e0, 0));
                for (int i0 = 0; i0 < b; i0 += 8) {
                    k_pull_spans<true, true><<<148, 64, 0, st[0]>>>(dm3 + (size_t)i0 * h * w, hmap + i0 * vs0, dp + i0 * vs0, 8, h, w, planes / 2, vs0, vs1, vs2, vs3, vs4);
                    if (mode & 1) CK(cudaMemcpyAsync(dm3 + (size_t)i0 * h * w, hm3 + (size_t)i0 * h * w, (size_t)8 * h * w, cudaMemcpyHostToDevice, st[1]));
                }
                cudaEvent_t ej;
                CK(cudaEventCreate(&ej)); CK(cudaEventRecord(ej, st[1])); CK(cudaStreamWaitEvent(st[0], ej, 0));
                CK(cudaEventRecord(e1, st[0]));
                CK(cudaEventSynchronize(e1));
                CK(cudaGetLastError());
                float ms;
                CK(cudaEventElapsedTime(&ms, e0, e1));
                const double mb = (double)b * planes * box * box * 4 / 1e6;
                if (rep) printf("pull kernel%s%s: %.3f ms  %.1f GB/s (spans only)\n", mode & 1 ? " + byte-mask copies (9.8 MB) on the copy engine" : "",
                                mode & 2 ? " + 8 host threads streaming host memory" : "", ms, mb / ms);
            }
            stop.store(true);
            for (auto &t : pool) t.join();
        }
    }
    // 128-byte aligned windows: x0 down to a multiple of 32 floats, width up to a multiple of 32
    for (int rep = 0; rep < 2; ++rep) {
        CK(cudaDeviceSynchronize());
        CK(cudaEventRecord(e0, st[0]));
        double mb = 0;
        for (int i = 0; i < b; ++i) {
            const int xa = (37 * i) % (w - box), y0 = (53 * i) % (h - box);
            const int x0 = xa & ~31, x1 = ((xa + box + 31) & ~31) > w ? w : ((xa + box + 31) & ~31);
            cudaMemcpy3DParms p = {};
            p.srcPtr = make_cudaPitchedPtr(hp + (size_t)i * planes * h * w, (size_t)w * 4, w, h);
            p.dstPtr = make_cudaPitchedPtr(dp + (size_t)i * planes * h * w, (size_t)w * 4, w, h);
            p.srcPos = make_cudaPos((size_t)x0 * 4, y0, 0);
            p.dstPos = p.srcPos;
            p.extent = make_cudaExtent((size_t)(x1 - x0) * 4, box, planes);
            p.kind = cudaMemcpyHostToDevice;
            CK(cudaMemcpy3DAsync(&p, st[0]));
            mb += (double)(x1 - x0) * 4 * box * planes / 1e6;
        }
        CK(cudaEventRecord(e1, st[0]));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("128-byte aligned windows: %.3f ms  %.1f MB  %.1f GB/s\n", ms, mb, mb / ms);
    }
    CK(cudaMemcpyAsync(dp, hp, n * 4, cudaMemcpyHostToDevice, st[0]));
    CK(cudaEventRecord(e0, st[0]));
    CK(cudaMemcpyAsync(dp, hp, n * 4, cudaMemcpyHostToDevice, st[0]));
    CK(cudaEventRecord(e1, st[0]));
    CK(cudaEventSynchronize(e1));
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    printf("whole tensor: %.3f ms  %.1f MB  %.1f GB/s\n", ms, n * 4 / 1e6, n * 4 / 1e6 / ms);
    return 0;
}
