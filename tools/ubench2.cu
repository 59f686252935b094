// ubench2.cu -- the K3 hot loop in isolation (no staging, no deferred passes, no barriers): what is the speed of light
// of the loop body on B200, packed (FFMA2) vs scalar (FFMA), at several occupancies?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench2 tools/ubench2.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk(float lo, float hi) { f32x2 d; asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(lo), "f"(hi)); return d; }
__device__ __forceinline__ void unpk(f32x2 v, float &lo, float &hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f32x2 ffma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }

constexpr int PX = 256;
// MODE 0: packed half-plane loop (as k_count); 1: scalar half-plane loop; 2: packed, FFMA2 only (no classification);
// 3: packed, min + sign count only (no band tracking); 4: band by FSETP chain instead of FMNMX3;
// 5: sign count by PRMT (sign-replicated bytes) + IDP.4A instead of 2 LEA.HI; 6: as 0 with the FMNMX on the integer pipe
// (VIMNMX on the float bits is NOT equivalent for negative values -- timing probe only)
template <int H, int THREADS, int MODE>
__global__ void __launch_bounds__(THREADS) k_loop(const float *consts, int reps, int *out)
{
    __shared__ __align__(16) float sP[12 * PX / 2];
    for (int i = threadIdx.x; i < 12 * PX / 2; i += THREADS) sP[i] = consts[i];
    __syncthreads();
    float hx[H], hy[H], T[H];
    int neg[H];
    uint32_t dmask[H];
#pragma unroll
    for (int k = 0; k < H; ++k) { hx[k] = 3.f + threadIdx.x + k; hy[k] = 7.f - k; T[k] = 1e-3f * (k + 1); neg[k] = 0; dmask[k] = 0; }
    for (int r = 0; r < reps; ++r) {
        uint32_t bit = 1u;
        const ulonglong2 *pp = reinterpret_cast<const ulonglong2 *>(sP);
        for (int g = 0; g < PX / 8; ++g, pp += 12, bit <<= 1) {
            float mabs[H];
#pragma unroll
            for (int k = 0; k < H; ++k) mabs[k] = __int_as_float(0x7f800000);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const ulonglong2 v0 = pp[3 * u], v1 = pp[3 * u + 1], v2 = pp[3 * u + 2];
#pragma unroll
                for (int k = 0; k < H; ++k) {
                    float ep0, ep1, em0, em1;
                    if (MODE == 1) {
                        float a0, a1, b0, b1, c0, c1, d0, d1, e0, e1, f0, f1;
                        unpk(v0.x, a0, a1); unpk(v0.y, b0, b1); unpk(v1.x, c0, c1);
                        unpk(v1.y, d0, d1); unpk(v2.x, e0, e1); unpk(v2.y, f0, f1);
                        ep0 = fmaf(hx[k], a0, fmaf(hy[k], b0, c0)); ep1 = fmaf(hx[k], a1, fmaf(hy[k], b1, c1));
                        em0 = fmaf(hx[k], d0, fmaf(hy[k], e0, f0)); em1 = fmaf(hx[k], d1, fmaf(hy[k], e1, f1));
                    } else {
                        const f32x2 hxx = pk(hx[k], hx[k]), hyy = pk(hy[k], hy[k]);
                        unpk(ffma2(hxx, v0.x, ffma2(hyy, v0.y, v1.x)), ep0, ep1);
                        unpk(ffma2(hxx, v1.y, ffma2(hyy, v2.x, v2.y)), em0, em1);
                    }
                    if (MODE == 4) {
                        const float m0 = fminf(ep0, em0), m1 = fminf(ep1, em1);
                        neg[k] += (int)(__float_as_uint(m0) >> 31) + (int)(__float_as_uint(m1) >> 31);
                        if (!(fabsf(m0) > T[k]) || !(fabsf(m1) > T[k])) mabs[k] = 0.f;
                    } else if (MODE == 5) {
                        const float m0 = fminf(ep0, em0), m1 = fminf(ep1, em1);
                        const unsigned sb = __byte_perm(__float_as_uint(m0), __float_as_uint(m1), 0xFFBB);
                        neg[k] = __dp4a((int)sb, 0x00010001, neg[k]);
                        mabs[k] = fminf(fminf(mabs[k], fabsf(m0)), fabsf(m1));
                    } else if (MODE == 2) {
                        neg[k] += __float_as_int(ep0) ^ __float_as_int(em1);  // cheap sink: 1 LOP3 + ... per 2 tests
                        hx[k] += ep1 * 0.f + em0 * 0.f;                        // keep all four live (2 FFMA)
                    } else {
                        const float m0 = fminf(ep0, em0), m1 = fminf(ep1, em1);
                        neg[k] += (int)(__float_as_uint(m0) >> 31) + (int)(__float_as_uint(m1) >> 31);
                        if (MODE != 3) mabs[k] = fminf(fminf(mabs[k], fabsf(m0)), fabsf(m1));
                    }
                }
            }
            if (MODE != 3 && MODE != 2) {
#pragma unroll
                for (int k = 0; k < H; ++k)
                    if (!(mabs[k] > T[k])) dmask[k] |= bit;
            }
        }
#pragma unroll
        for (int k = 0; k < H; ++k) hy[k] += 1e-3f;  // perturb so that repetitions cannot be merged
    }
    int s = 0;
#pragma unroll
    for (int k = 0; k < H; ++k) s += neg[k] + (int)dmask[k] + (int)hx[k];
    if (s == 0x7fffffff) out[0] = s;
}

template <int H, int THREADS, int MODE>
static void run(const char *name, int ctas_per_sm, const float *dc, int *out)
{
    int sms = 148, khz = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const int reps = 400;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k_loop<H, THREADS, MODE><<<sms * ctas_per_sm, THREADS>>>(dc, 10, out);
    float best = 1e30f;
    for (int i = 0; i < 3; ++i) {
        cudaEventRecord(e0);
        k_loop<H, THREADS, MODE><<<sms * ctas_per_sm, THREADS>>>(dc, reps, out);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    const double tests = (double)sms * ctas_per_sm * THREADS * H * PX * reps;
    const double warp_tests_per_smsp = tests / 32.0 / (sms * 4.0);
    const double cyc = best * 1e-3 * khz * 1e3;
    int occ = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_loop<H, THREADS, MODE>, THREADS, 0);
    printf("%-34s H=%d T=%3d ctas/sm=%2d (max %2d): %7.3f ms  %6.2f T tests/s  %5.2f cyc/test/SMSP\n", name, H, THREADS,
           ctas_per_sm, occ, best, tests / best / 1e9, cyc / warp_tests_per_smsp);
}

int main()
{
    float h[12 * PX / 2];
    for (int i = 0; i < 12 * PX / 2; ++i) h[i] = (float)((i * 2654435761u) % 2001) / 1000.f - 1.f;
    float *dc; int *out;
    cudaMalloc(&dc, sizeof(h)); cudaMalloc(&out, 4);
    cudaMemcpy(dc, h, sizeof(h), cudaMemcpyHostToDevice);
    run<4, 128, 0>("packed full", 4, dc, out);
    run<4, 128, 0>("packed full", 8, dc, out);
    run<4, 128, 1>("scalar full", 4, dc, out);
    run<4, 128, 1>("scalar full", 8, dc, out);
    run<4, 128, 2>("packed FFMA2 only", 8, dc, out);
    run<4, 128, 3>("packed min+count", 8, dc, out);
    run<8, 64, 0>("packed full", 8, dc, out);
    run<8, 64, 0>("packed full", 12, dc, out);
    run<2, 128, 0>("packed full", 8, dc, out);
    run<2, 128, 0>("packed full", 12, dc, out);
    run<4, 64, 0>("packed full", 16, dc, out);
    run<4, 32, 0>("packed full", 32, dc, out);
    run<4, 128, 4>("packed, FSETP band", 8, dc, out);
    run<4, 128, 5>("packed, PRMT+IDP4A count", 8, dc, out);
    run<4, 128, 0>("packed full", 5, dc, out);
    run<4, 128, 0>("packed full", 6, dc, out);
    run<4, 128, 0>("packed full", 10, dc, out);
    run<4, 128, 0>("packed full", 12, dc, out);
    return 0;
}
