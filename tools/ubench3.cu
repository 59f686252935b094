// ubench3.cu -- per-instruction issue cost on sm_100a with REGISTER operands only (no immediates, no constant bank,
// no uniform registers): 8 independent chains per thread, 8 warps per SM sub-partition.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench3 tools/ubench3.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;

#define A_FFMA(i)   asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(x[i]) : "f"(y[i]), "f"(z[i]))
#define A_FFMA2(i)  asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(p[i]) : "l"(q[i]), "l"(r[i]))
#define A_FFMA2B(i) asm volatile("{ .reg .b64 t; mov.b64 t, {%1, %1}; fma.rn.f32x2 %0, t, %2, %0; }" : "+l"(p[i]) : "f"(y[i]), "l"(r[i]))
#define A_FADD(i)   asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(x[i]) : "f"(y[i]))
#define A_FMNMX(i)  asm volatile("min.f32 %0, %0, %1;" : "+f"(w[i]) : "f"(z[i]))
#define A_FMNMX3(i) asm volatile("min.f32 %0, %0, %1, %2;" : "+f"(w[i]) : "f"(z[i]), "f"(y[i]))
#define A_LEAHI(i)  asm volatile("{ .reg .b32 t; shr.u32 t, %1, 31; add.s32 %0, %0, t; }" : "+r"(n[i]) : "r"(m[i]))
#define A_LOP3(i)   asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(n[i]) : "r"(m[i]), "r"(o[i]))
#define A_IADD3(i)  asm volatile("{ .reg .b32 t; add.s32 t, %1, %2; add.s32 %0, %0, t; }" : "+r"(n[i]) : "r"(m[i]), "r"(o[i]))
#define A_PRMT(i)   asm volatile("prmt.b32 %0, %0, %1, 0x7733;" : "+r"(n[i]) : "r"(m[i]))
#define A_IDP(i)    asm volatile("dp4a.s32.s32 %0, %1, %2, %0;" : "+r"(n[i]) : "r"(m[i]), "r"(o[i]))
#define A_FSETPSEL(i) asm volatile("{ .reg .pred q; setp.gt.f32 q, %1, %2; selp.b32 %0, %3, %0, q; }" : "+r"(n[i]) : "f"(z[i]), "f"(y[i]), "r"(m[i]))
#define A_IMAD(i)   asm volatile("mad.lo.s32 %0, %1, %2, %0;" : "+r"(n[i]) : "r"(m[i]), "r"(o[i]))

#define KERNEL(NAME, BODY)                                                                      \
    __global__ void __launch_bounds__(256) NAME(float *out, int iters)                          \
    {                                                                                           \
        float x[8], y[8], z[8], w[8];                                                           \
        u64 p[8], q[8], r[8];                                                                   \
        int n[8], m[8], o[8];                                                                   \
        _Pragma("unroll") for (int i = 0; i < 8; ++i) {                                         \
            x[i] = threadIdx.x * 0.001f + i; y[i] = 1.0f + 1e-7f * (threadIdx.x + i);           \
            z[i] = 1e-9f * (threadIdx.x + 3 * i + 1); w[i] = 5.f + i + threadIdx.x;             \
            n[i] = threadIdx.x * 7 + i; m[i] = threadIdx.x * 13 + 5 * i + 1; o[i] = threadIdx.x * 3 + i + 11; \
            asm volatile("mov.b64 %0, {%1, %2};" : "=l"(p[i]) : "f"(x[i]), "f"(w[i]));          \
            asm volatile("mov.b64 %0, {%1, %2};" : "=l"(q[i]) : "f"(y[i]), "f"(y[i]));          \
            asm volatile("mov.b64 %0, {%1, %2};" : "=l"(r[i]) : "f"(z[i]), "f"(z[i]));          \
        }                                                                                       \
        for (int it = 0; it < iters; ++it) {                                                    \
            _Pragma("unroll") for (int i = 0; i < 8; ++i) { BODY }                              \
        }                                                                                       \
        float s = 0.f;                                                                          \
        _Pragma("unroll") for (int i = 0; i < 8; ++i) {                                         \
            float lo, hi;                                                                       \
            asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p[i]));             \
            s += x[i] + w[i] + lo + hi + (float)n[i];                                           \
        }                                                                                       \
        if (s == 123.456f) out[0] = s;                                                          \
    }

KERNEL(k_ffma, A_FFMA(i);)
KERNEL(k_ffma2, A_FFMA2(i);)
KERNEL(k_ffma2b, A_FFMA2B(i);)
KERNEL(k_fadd, A_FADD(i);)
KERNEL(k_fmnmx, A_FMNMX(i);)
KERNEL(k_fmnmx3, A_FMNMX3(i);)
KERNEL(k_leahi, A_LEAHI(i);)
KERNEL(k_lop3, A_LOP3(i);)
KERNEL(k_iadd3, A_IADD3(i);)
KERNEL(k_prmt, A_PRMT(i);)
KERNEL(k_idp, A_IDP(i);)
KERNEL(k_fsetpsel, A_FSETPSEL(i);)
KERNEL(k_imad, A_IMAD(i);)
KERNEL(k_ffma_fmnmx, A_FFMA(i); A_FMNMX(i);)
KERNEL(k_2ffma_fmnmx, A_FFMA(i); A_FADD(i); A_FMNMX(i);)
KERNEL(k_ffma2_fmnmx, A_FFMA2(i); A_FMNMX(i);)
KERNEL(k_ffma2_2fmnmx, A_FFMA2(i); A_FMNMX(i); A_LEAHI(i);)
KERNEL(k_ffma2_lea, A_FFMA2(i); A_LEAHI(i);)
KERNEL(k_fmnmx_lea, A_FMNMX(i); A_LEAHI(i);)
KERNEL(k_ffma2_idp, A_FFMA2(i); A_IDP(i);)
KERNEL(k_ffma2_imad, A_FFMA2(i); A_IMAD(i);)
KERNEL(k_body, A_FFMA2(i); A_FFMA2B(i); A_FMNMX(i); A_LEAHI(i); if (i & 1) { A_FMNMX3(i); })

template <typename K>
static void run(const char *name, K kern, int body)
{
    int sms = 148, khz = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    float *out;
    cudaMalloc(&out, 4);
    const int iters = 1 << 14, blocks = sms * 4;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    kern<<<blocks, 256>>>(out, 64);
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0);
        kern<<<blocks, 256>>>(out, iters);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    const double cyc = best * 1e-3 * khz * 1e3 / ((double)iters * 8.0 * 8.0);  // per unrolled step (one of 8) per warp
    printf("%-28s %7.3f ms  %6.2f cycles per step (%d instr -> %.2f cyc/instr)\n", name, best, cyc, body, cyc / body);
    cudaFree(out);
}

int main()
{
    run("FFMA r,r,r", k_ffma, 1);
    run("FFMA2 r2,r2,r2", k_ffma2, 1);
    run("FFMA2 bcast.F32,r2,r2", k_ffma2b, 1);
    run("FADD", k_fadd, 1);
    run("FMNMX", k_fmnmx, 1);
    run("FMNMX3", k_fmnmx3, 1);
    run("LEA.HI (shr+add)", k_leahi, 1);
    run("LOP3", k_lop3, 1);
    run("IADD3", k_iadd3, 1);
    run("PRMT", k_prmt, 1);
    run("IDP.4A", k_idp, 1);
    run("FSETP+SEL", k_fsetpsel, 2);
    run("IMAD", k_imad, 1);
    run("FFMA + FMNMX", k_ffma_fmnmx, 2);
    run("FFMA + FADD + FMNMX", k_2ffma_fmnmx, 3);
    run("FFMA2 + FMNMX", k_ffma2_fmnmx, 2);
    run("FFMA2 + FMNMX + LEA.HI", k_ffma2_2fmnmx, 3);
    run("FFMA2 + LEA.HI", k_ffma2_lea, 2);
    run("FMNMX + LEA.HI", k_fmnmx_lea, 2);
    run("FFMA2 + IDP.4A", k_ffma2_idp, 2);
    run("FFMA2 + IMAD", k_ffma2_imad, 2);
    run("2 FFMA2 + FMNMX + LEA + .5 FMNMX3", k_body, 5);
    return 0;
}
