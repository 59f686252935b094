// ubench.cu -- instruction-mix microbenchmarks for sm_100a (B200): which pipe / issue budget does each op of the
// K3 inner loop cost?  Standalone:  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench tools/ubench.cu
// Prints, per mix, the cycles one SM sub-partition (SMSP) needs per loop body (8 warps resident per SMSP).
#include <cstdio>
#include <cuda_runtime.h>

typedef unsigned long long u64;
#define PK(d, lo, hi) asm volatile("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(lo), "f"(hi))

template <int MIX>
__global__ void __launch_bounds__(256) k_mix(float *out, int iters, float fa, float fb, int ia)
{
    float x[8], y[8];
    u64 p[8], q, r;
    int n[8];
    PK(q, fa, fa);
    PK(r, fb, fb);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        x[i] = threadIdx.x + i; y[i] = fa * i; n[i] = ia + i + (int)threadIdx.x * 7;
        PK(p[i], x[i], y[i]);
    }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MIX == 0) {  // 8 FFMA
                asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(fa), "f"(fb));
            } else if (MIX == 1) {  // 8 FFMA2
                asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(q), "l"(r));
            } else if (MIX == 2) {  // 8 FFMA2 + 8 LOP3 (ALU)
                asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(q), "l"(r));
                asm volatile("xor.b32 %0, %0, %1;" : "+r"(n[i]) : "r"(ia));
            } else if (MIX == 3) {  // 8 FFMA2 + 16 LOP3
                asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(q), "l"(r));
                asm volatile("xor.b32 %0, %0, %1;" : "+r"(n[i]) : "r"(ia));
                asm volatile("xor.b32 %0, %0, %1;" : "+r"(n[i]) : "r"(it));
            } else if (MIX == 4) {  // 16 FFMA + 8 LOP3
                asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(fa), "f"(fb));
                asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(y[i]) : "f"(fa), "f"(fb));
                asm volatile("xor.b32 %0, %0, %1;" : "+r"(n[i]) : "r"(ia));
            } else if (MIX == 5) {  // 16 FFMA + 16 LOP3
                asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(fa), "f"(fb));
                asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(y[i]) : "f"(fa), "f"(fb));
                asm volatile("xor.b32 %0, %0, %1;" : "+r"(n[i]) : "r"(ia));
                asm volatile("xor.b32 %0, %0, %1;" : "+r"(n[i]) : "r"(it));
            } else if (MIX == 6) {  // 8 FFMA2 + 8 FADD
                asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(q), "l"(r));
                asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(x[i]) : "f"(fb));
            } else if (MIX == 7) {  // 8 LOP3 only
                asm volatile("xor.b32 %0, %0, %1;" : "+r"(n[i]) : "r"(it));
            } else if (MIX == 8) {  // 8 FMNMX3
                asm volatile("min.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(fa), "f"(y[i]));
            } else if (MIX == 9) {  // 8 FADD2
                asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(q));
            } else if (MIX == 10) {  // 8 FFMA (3 distinct regs) + 8 FADD + 8 LOP3 + 8 FSETP-ish (set.gt)
                asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(fa), "f"(y[i]));
                asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(y[i]) : "f"(fb));
                asm volatile("xor.b32 %0, %0, %1;" : "+r"(n[i]) : "r"(ia));
            } else if (MIX == 11) {  // 8 FFMA2 with a 32-bit broadcast-able operand {fa,fa} built in-loop-invariant
                asm volatile("fma.rn.f32x2 %0, %1, %0, %2;" : "+l"(p[i]) : "l"(q), "l"(p[(i + 1) & 7]));
            } else if (MIX == 13) {  // K3 body for 2 tests x 4: 2 FFMA2 + 2 FADD + 2 FADD + 2 LOP3(LEA.HI) + 2 FSETP(min.f32)
                asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(q), "l"(r));
                asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(x[i]) : "f"(fb));
                asm volatile("xor.b32 %0, %0, %1;" : "+r"(n[i]) : "r"(ia));
                if (i & 1) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(y[i]) : "f"(fb));
                else asm volatile("min.f32 %0, %0, %1;" : "+f"(y[i]) : "f"(fb));
            } else if (MIX == 14) {  // scalar K3 body per test x 2: 4 FFMA + 2 FADD + 1 LOP3 + 1 FMNMX, x2
                asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(fa), "f"(fb));
                if (i & 1) asm volatile("xor.b32 %0, %0, %1;" : "+r"(n[i]) : "r"(ia));
                else if (i & 2) asm volatile("min.f32 %0, %0, %1;" : "+f"(y[i]) : "f"(fb));
                if (i < 4) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(y[i + 4]) : "f"(fb));
            } else if (MIX == 15) {  // 8 FFMA2 + 8 FMNMX (ALU, fp)
                asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(q), "l"(r));
                asm volatile("min.f32 %0, %0, %1;" : "+f"(y[i]) : "f"(fb));
            } else if (MIX == 16) {  // 8 FMNMX
                asm volatile("min.f32 %0, %0, %1;" : "+f"(y[i]) : "f"(fb));
            } else if (MIX == 17) {  // 8 FFMA + 8 FMNMX
                asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(fa), "f"(fb));
                asm volatile("min.f32 %0, %0, %1;" : "+f"(y[i]) : "f"(fb));
            } else if (MIX == 18) {  // 16 FFMA + 8 FMNMX
                asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(fa), "f"(fb));
                asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(fb), "f"(fa));
                asm volatile("min.f32 %0, %0, %1;" : "+f"(y[i]) : "f"(fb));
            } else if (MIX == 12) {  // 4 FFMA2 + 4 FADD + 4 LOP3 + 4 FSETP (the K3 body shape per 2 tests... x4)
                asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(q), "l"(r));
                if (i & 1) {
                    asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(x[i]) : "f"(fb));
                    asm volatile("xor.b32 %0, %0, %1;" : "+r"(n[i]) : "r"(ia));
                }
            }
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        float lo, hi;
        asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p[i]));
        s += x[i] + y[i] + lo + hi + (float)n[i];
    }
    if (s == 123.456f) out[0] = s;
}

template <int MIX>
static void run(const char *name, int body_instr)
{
    int dev = 0, sms = 148, khz = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
    float *out;
    cudaMalloc(&out, 4);
    const int iters = 1 << 15, blocks = sms * 4;  // 4 x 256 threads per SM = 8 warps per SMSP
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    k_mix<MIX><<<blocks, 256>>>(out, 256, 1.0000001f, 1e-9f, 3);
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0);
        k_mix<MIX><<<blocks, 256>>>(out, iters, 1.0000001f, 1e-9f, 3);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    // cycles per loop body per SMSP at the nominal max clock: 8 warps per SMSP each run `iters` bodies
    const double cyc = best * 1e-3 * (double)khz * 1e3 / ((double)iters * 8.0);
    printf("%-44s %8.3f ms   %6.2f cyc/body/warp @%d MHz  (%d instr/body -> %.2f cyc/instr)\n", name, best, cyc, khz / 1000,
           body_instr, cyc / body_instr);
    cudaFree(out);
}

int main()
{
    run<0>("8 FFMA", 8);
    run<1>("8 FFMA2", 8);
    run<2>("8 FFMA2 + 8 LOP3", 16);
    run<3>("8 FFMA2 + 16 LOP3", 24);
    run<4>("16 FFMA + 8 LOP3", 24);
    run<5>("16 FFMA + 16 LOP3", 32);
    run<6>("8 FFMA2 + 8 FADD", 16);
    run<7>("8 LOP3", 8);
    run<8>("8 FMNMX3", 8);
    run<9>("8 FADD2", 8);
    run<10>("8 FFMA(3reg) + 8 FADD + 8 LOP3", 24);
    run<11>("8 FFMA2 (3 x 64-bit regs)", 8);
    run<12>("8 FFMA2 + 4 FADD + 4 LOP3", 16);
    run<13>("8 FFMA2 + 12 FADD + 8 LOP3 + 4 FMNMX", 32);
    run<14>("8 FFMA + 4 FADD + 4 LOP3 + 2 FMNMX", 18);
    run<15>("8 FFMA2 + 8 FMNMX", 16);
    run<16>("8 FMNMX", 8);
    run<17>("8 FFMA + 8 FMNMX", 16);
    run<18>("16 FFMA + 8 FMNMX", 24);
    return 0;
}
