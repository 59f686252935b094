// Probe (GPU box): can a kernel that needs the SM's largest shared-memory carve-out (224 KB, like the tensor-pipe counting
// kernel) start on an SM on which a small no-shared-memory CTA (like the span-pull kernel) is already resident?
// build: nvcc -O2 -gencode arch=compute_100a,code=sm_100a tools/coresidency_probe.cu -o tools/_build/coresidency_probe
#include <cuda_runtime.h>
#include <stdio.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)
__global__ void __launch_bounds__(64, 32) k_small(long long cycles, int *sink)
{
    const long long t0 = clock64();
    while (clock64() - t0 < cycles) { }
    if (cycles < 0) *sink = 1;
}
__global__ void __launch_bounds__(768, 1) k_big(int *sink)
{
    extern __shared__ unsigned char sm[];
    sm[threadIdx.x] = 1;
    __syncthreads();
    if (sm[0] == 7) *sink = 2;
}
int main()
{
    int *sink;
    CK(cudaMalloc(&sink, 4));
    const int big_smem = 228608;
    CK(cudaFuncSetAttribute(k_big, cudaFuncAttributeMaxDynamicSharedMemorySize, big_smem));
    cudaStream_t a, b;
    CK(cudaStreamCreateWithFlags(&a, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&b, cudaStreamNonBlocking));
    cudaEvent_t e0, ea, eb;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&ea)); CK(cudaEventCreate(&eb));
    for (int mode = 0; mode < 4; ++mode) {
        // mode 0: small kernel as it is; 1: carve-out hint MaxShared; 2: hint + 1 KB dynamic smem; 3: small kernel launched second
        if (mode >= 1) CK(cudaFuncSetAttribute(k_small, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        for (int rep = 0; rep < 3; ++rep) {
            CK(cudaDeviceSynchronize());
            CK(cudaEventRecord(e0, a));
            CK(cudaStreamWaitEvent(b, e0, 0));
            if (mode == 3) {
                k_big<<<148, 768, big_smem, b>>>(sink);
                k_small<<<148, 64, 0, a>>>(400000, sink);
            } else {
                k_small<<<148, 64, mode == 2 ? 1024 : 0, a>>>(400000, sink);   // ~200 us
                k_big<<<148, 768, big_smem, b>>>(sink);
            }
            CK(cudaEventRecord(ea, a));
            CK(cudaEventRecord(eb, b));
            CK(cudaDeviceSynchronize());
            CK(cudaGetLastError());
            float ta, tb;
            CK(cudaEventElapsedTime(&ta, e0, ea));
            CK(cudaEventElapsedTime(&tb, e0, eb));
            if (rep) printf("mode %d: small kernel done at %.1f us, 224 KB kernel done at %.1f us  -> %s\n", mode, ta * 1e3, tb * 1e3,
                            tb < ta * 0.8f ? "co-resident" : "waited");
        }
    }
    return 0;
}
