// pvv_host_simd.cpp -- host-side helpers of the host-buffer entry points (pvv_host.cu): narrowing of wide masks to one
// byte per pixel and the foreground bounding box of a run of rows.  Plain C++ (compiled by the host compiler, no CUDA): the
// AVX-512 bodies are picked at run time (__builtin_cpu_supports), everything has a portable scalar body.
//
// Measured on the 16-thread Xeon of a B200 box (tools/narrow_probe.cpp, 78.6 MB int64 mask, 8 threads): scalar 65 GB/s,
// AVX-512 (vpmovqb) 104 GB/s, AVX-512 + software prefetch 132 GB/s.
#include <immintrin.h>
#include <stdint.h>
#include <string.h>

#include "pvv_host_simd.h"

namespace {

// ---- int64 -> byte ------------------------------------------------------------------------------------------------
void narrow_i64_scalar(const uint64_t *w, uint8_t *dst, size_t n, int eq1)
{
    size_t i = 0;
    if (eq1) {
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

#if defined(__x86_64__)
constexpr int kPrefetchAhead = 16 * 64;   // elements (8 KB): measured best of 0 / 4 / 16 blocks of 64 pixels

__attribute__((target("avx512f,avx512bw,avx512vl"))) void narrow_i64_avx512(const uint64_t *w, uint8_t *dst, size_t n, int eq1)
{
    size_t i = 0;
    if (eq1) {
        const __m512i one = _mm512_set1_epi64(1);
        for (; i + 64 <= n; i += 64) {
            for (int k = 0; k < 8; ++k) _mm_prefetch((const char *)(w + i + kPrefetchAhead + 8 * k), _MM_HINT_T0);
            uint64_t m = 0;
            for (int k = 0; k < 8; ++k)
                m |= (uint64_t)_mm512_cmpeq_epi64_mask(_mm512_loadu_si512(w + i + 8 * k), one) << (8 * k);
            _mm512_storeu_si512(dst + i, _mm512_maskz_set1_epi8((__mmask64)m, 1));
        }
    } else {
        for (; i + 64 <= n; i += 64) {
            for (int k = 0; k < 8; ++k) _mm_prefetch((const char *)(w + i + kPrefetchAhead + 8 * k), _MM_HINT_T0);
            __m128i a[8];
            for (int k = 0; k < 8; ++k) a[k] = _mm512_cvtepi64_epi8(_mm512_loadu_si512(w + i + 8 * k));   // vpmovqb
            for (int k = 0; k < 8; k += 2) _mm_storeu_si128((__m128i *)(dst + i + 8 * k), _mm_unpacklo_epi64(a[k], a[k + 1]));
        }
    }
    narrow_i64_scalar(w + i, dst + i, n - i, eq1);
}

__attribute__((target("avx512f,avx512bw,avx512vl"))) void narrow_i32_avx512(const uint32_t *w, uint8_t *dst, size_t n, int eq1)
{
    size_t i = 0;
    const __m512i one = _mm512_set1_epi32(1);
    for (; i + 64 <= n; i += 64) {
        for (int k = 0; k < 4; ++k) _mm_prefetch((const char *)(w + i + 2 * kPrefetchAhead + 16 * k), _MM_HINT_T0);
        if (eq1) {
            uint64_t m = 0;
            for (int k = 0; k < 4; ++k)
                m |= (uint64_t)_mm512_cmpeq_epi32_mask(_mm512_loadu_si512(w + i + 16 * k), one) << (16 * k);
            _mm512_storeu_si512(dst + i, _mm512_maskz_set1_epi8((__mmask64)m, 1));
        } else {
            for (int k = 0; k < 4; ++k)
                _mm_storeu_si128((__m128i *)(dst + i + 16 * k), _mm512_cvtepi32_epi8(_mm512_loadu_si512(w + i + 16 * k)));  // vpmovdb
        }
    }
    for (; i < n; ++i) dst[i] = eq1 ? (w[i] == 1u ? 1 : 0) : (uint8_t)w[i];
}

bool have_avx512()
{
    static const bool ok = __builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512bw") && __builtin_cpu_supports("avx512vl");
    return ok;
}
#else
bool have_avx512() { return false; }
#endif

}  // namespace

void pvv_narrow_i64(const int64_t *src, uint8_t *dst, size_t n, int eq1)
{
#if defined(__x86_64__)
    if (have_avx512()) { narrow_i64_avx512((const uint64_t *)src, dst, n, eq1); return; }
#endif
    narrow_i64_scalar((const uint64_t *)src, dst, n, eq1);
}

void pvv_narrow_i32(const int32_t *src, uint8_t *dst, size_t n, int eq1)
{
#if defined(__x86_64__)
    if (have_avx512()) { narrow_i32_avx512((const uint32_t *)src, dst, n, eq1); return; }
#endif
    for (size_t i = 0; i < n; ++i) dst[i] = eq1 ? (src[i] == 1 ? 1 : 0) : (uint8_t)(uint32_t)src[i];
}

// Write the cache lines of [p, p + n) back to memory (they stay valid in the cache).  See pvv_host.cu: a copy engine that
// reads bytes a core has just written, after that core has gone to sleep, crawls.
#if defined(__x86_64__)
__attribute__((target("clwb"))) static void writeback_clwb(const uint8_t *p, size_t n)
{
    const uintptr_t a = (uintptr_t)p & ~(uintptr_t)63, e = (uintptr_t)p + n;
    for (uintptr_t q = a; q < e; q += 64) _mm_clwb((void *)q);
    _mm_sfence();
}
static bool have_clwb()
{
    static const bool ok = [] {
        unsigned a = 0, b = 0, c = 0, d = 0;
        __asm__ volatile("cpuid" : "=a"(a), "=b"(b), "=c"(c), "=d"(d) : "a"(7), "c"(0));
        return (b >> 24) & 1u;
    }();
    return ok;
}
#endif
void pvv_writeback(const uint8_t *p, size_t n)
{
#if defined(__x86_64__)
    if (have_clwb()) writeback_clwb(p, n);
#else
    (void)p; (void)n;
#endif
}

void pvv_narrow_f32(const float *src, uint8_t *dst, size_t n, int eq1)
{
    if (eq1) { for (size_t i = 0; i < n; ++i) dst[i] = src[i] == 1.f ? 1 : 0; }
    else { for (size_t i = 0; i < n; ++i) dst[i] = (uint8_t)(long long)src[i]; }   // truncation toward zero, like .byte()
}

// Foreground (non-zero byte) statistics of `rows` rows of `w` bytes: returns the number of foreground pixels and adds to
// *span_px the summed row extents (last - first + 1 of every row that has foreground) -- what a row-span transfer of the
// image moves per plane.
long long pvv_fg_extent(const uint8_t *bytes, int w, int rows, long long *span_px)
{
    long long count = 0, span = 0;
    for (int r = 0; r < rows; ++r) {
        const uint8_t *row = bytes + (size_t)r * w;
        int first = -1, last = -1, x = 0;
        long long c = 0;
        for (; x + 8 <= w; x += 8) {
            uint64_t v;
            memcpy(&v, row + x, 8);
            if (!v) continue;
            // bit 7 of every non-zero byte
            const uint64_t nz = (((v & 0x7f7f7f7f7f7f7f7full) + 0x7f7f7f7f7f7f7f7full) | v) & 0x8080808080808080ull;
            c += __builtin_popcountll(nz);
            if (first < 0) first = x + (__builtin_ctzll(nz) >> 3);
            last = x + 7 - (__builtin_clzll(nz) >> 3);
        }
        for (; x < w; ++x)
            if (row[x]) { ++c; if (first < 0) first = x; last = x; }
        if (c) { count += c; span += last - first + 1; }
    }
    *span_px += span;
    return count;
}
