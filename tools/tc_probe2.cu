// tc_probe2.cu -- tcgen05.ld (TMEM -> registers) latency and throughput on one SM of a B200, as seen by the voting warps
// of the tensor-pipe K3: each warp repeatedly reads 64 columns of its lane quarter (2 x LDTM.x32) and waits for them.
// Prints cycles per (2 x LDTM.x32 + wait) for 1..16 warps per SM, and with 1 / 2 / 4 loads in flight per wait.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); return 1; } } while (0)

__device__ __forceinline__ void ld32(uint32_t taddr, uint32_t (&v)[32])
{
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
          "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
          "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr));
}

template <int DEPTH>
__global__ void __launch_bounds__(512) probe(int iters, long long *cycles, uint32_t *sink)
{
    __shared__ uint32_t s_tmem;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&s_tmem)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t base = s_tmem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)((warp >> 2) * 64);
    uint32_t acc = 0;
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
        uint32_t v[DEPTH][32];
#pragma unroll
        for (int q = 0; q < DEPTH; ++q) ld32(base + (uint32_t)((q & 1) * 32 + (i & 1) * 256), v[q]);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
        for (int q = 0; q < DEPTH; ++q) acc ^= v[q][0] ^ v[q][31];
    }
    const long long t1 = clock64();
    if ((threadIdx.x & 31) == 0) cycles[warp] = t1 - t0;
    sink[threadIdx.x] = acc;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(s_tmem), "r"(512));
}

int main()
{
    long long *dc;
    uint32_t *ds;
    CK(cudaMalloc(&dc, 16 * 8));
    CK(cudaMalloc(&ds, 512 * 4));
    const int iters = 2000;
    for (int depth : {1, 2, 4})
        for (int warps : {1, 4, 8, 16}) {
            if (depth == 1) probe<1><<<1, warps * 32>>>(iters, dc, ds);
            if (depth == 2) probe<2><<<1, warps * 32>>>(iters, dc, ds);
            if (depth == 4) probe<4><<<1, warps * 32>>>(iters, dc, ds);
            CK(cudaDeviceSynchronize());
            long long c[16];
            CK(cudaMemcpy(c, dc, sizeof(c), cudaMemcpyDeviceToHost));
            long long mx = 0;
            for (int w = 0; w < warps; ++w) mx = c[w] > mx ? c[w] : mx;
            printf("%2d warps, %d x LDTM.x32 per wait: %7.1f cycles per iteration  (%.1f cycles per LDTM.x32 per SM sub-partition)\n", warps,
                   depth, (double)mx / iters, (double)mx / iters / depth / ((warps + 3) / 4));
        }
    return 0;
}
