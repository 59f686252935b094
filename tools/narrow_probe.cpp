// Probe (GPU box host): how fast do N host threads narrow an int64 mask to bytes?  scalar / AVX-512 / AVX-512 + prefetch.
// build: g++ -O3 -std=c++17 -pthread tools/narrow_probe.cpp -o tools/_build/narrow_probe
#include <immintrin.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <atomic>
#include <chrono>
#include <thread>
#include <vector>

static void narrow_scalar(const uint64_t *w, uint8_t *dst, size_t n)
{
    for (size_t i = 0; i + 8 <= n; i += 8) {
        uint64_t o = 0;
        for (int k = 0; k < 8; ++k) o |= (w[i + k] & 0xffu) << (8 * k);
        memcpy(dst + i, &o, 8);
    }
}
__attribute__((target("avx512f,avx512bw,avx512vl"))) static void narrow_avx512(const uint64_t *w, uint8_t *dst, size_t n, int pf)
{
    for (size_t i = 0; i + 64 <= n; i += 64) {
        if (pf) for (int k = 0; k < 8; ++k) _mm_prefetch((const char *)(w + i + 64 * pf + 8 * k), _MM_HINT_T0);
        __m128i a[8];
        for (int k = 0; k < 8; ++k) a[k] = _mm512_cvtepi64_epi8(_mm512_loadu_si512(w + i + 8 * k));
        for (int k = 0; k < 8; k += 2) {
            const __m128i v = _mm_unpacklo_epi64(a[k], a[k + 1]);
            _mm_storeu_si128((__m128i *)(dst + i + 8 * k), v);
        }
    }
}
__attribute__((target("avx2"))) static void narrow_avx2(const uint64_t *w, uint8_t *dst, size_t n)
{
    // low bytes of 4 int64 per ymm: shuffle bytes 0 and 8 of each 128-bit lane to the lane's bottom, then gather the lanes
    const __m256i sh = _mm256_setr_epi8(0, 8, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, 0, 8, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1);
    for (size_t i = 0; i + 8 <= n; i += 8) {
        const __m256i x = _mm256_shuffle_epi8(_mm256_loadu_si256((const __m256i *)(w + i)), sh);
        const __m256i y = _mm256_shuffle_epi8(_mm256_loadu_si256((const __m256i *)(w + i + 4)), sh);
        const uint64_t o = (uint64_t)(uint16_t)_mm256_extract_epi16(x, 0) | (uint64_t)(uint16_t)_mm256_extract_epi16(x, 8) << 16 |
                           (uint64_t)(uint16_t)_mm256_extract_epi16(y, 0) << 32 | (uint64_t)(uint16_t)_mm256_extract_epi16(y, 8) << 48;
        memcpy(dst + i, &o, 8);
    }
}

int main()
{
    const size_t n = (size_t)32 * 480 * 640;
    const int sets = 3;
    std::vector<uint64_t *> src(sets);
    for (auto &p : src) { p = (uint64_t *)aligned_alloc(4096, n * 8); for (size_t i = 0; i < n; ++i) p[i] = (i * 2654435761u >> 7) & 1; }
    uint8_t *dst = (uint8_t *)aligned_alloc(4096, n), *ref = (uint8_t *)aligned_alloc(4096, n);
    narrow_scalar(src[0], ref, n);
    for (int variant = 0; variant < 5; ++variant) {
        memset(dst, 0xee, n);
        for (int threads : {4, 6, 8, 10, 12, 16}) {
            double best = 1e9;
            for (int rep = 0; rep < 6; ++rep) {
                const uint64_t *w = src[rep % sets];
                std::atomic<size_t> next(0);
                const size_t piece = 1 << 16;
                auto t0 = std::chrono::steady_clock::now();
                auto work = [&] {
                    for (;;) {
                        const size_t c = next.fetch_add(1) * piece;
                        if (c >= n) return;
                        const size_t m = c + piece <= n ? piece : n - c;
                        if (variant == 0) narrow_scalar(w + c, dst + c, m);
                        else if (variant == 1) narrow_avx512(w + c, dst + c, m, 0);
                        else if (variant == 2) narrow_avx512(w + c, dst + c, m, 4);
                        else if (variant == 3) narrow_avx512(w + c, dst + c, m, 16);
                        else narrow_avx2(w + c, dst + c, m);
                    }
                };
                std::vector<std::thread> pool;
                for (int t = 0; t < threads; ++t) pool.emplace_back(work);
                for (auto &t : pool) t.join();
                const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
                if (rep && ms < best) best = ms;
            }
            printf("variant %d (%s) threads %2d: %.3f ms  %.1f GB/s  ok=%d\n", variant,
                   variant == 0 ? "scalar" : variant == 1 ? "avx512" : variant == 2 ? "avx512+pf4" : variant == 3 ? "avx512+pf16" : "avx2", threads, best,
                   n * 8 / 1e6 / best, memcmp(dst, ref, n) == 0);
        }
    }
    return 0;
}
