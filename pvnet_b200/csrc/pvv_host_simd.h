// pvv_host_simd.h -- host-side helpers of the host-buffer entry points (pvv_host_simd.cpp; plain C++, no CUDA).
#pragma once
#include <stddef.h>
#include <stdint.h>

// internal to libpvvote.so (not part of include/pvvote.h); tests/test_host_simd.py builds the file on its own with
// -DPVV_HOST_SIMD_TEST to reach the functions through ctypes
#ifdef PVV_HOST_SIMD_TEST
#define PVV_HOST_LOCAL extern "C" __attribute__((visibility("default")))
#else
#define PVV_HOST_LOCAL __attribute__((visibility("hidden")))
#endif

// n mask elements -> one byte each: the value the foreground rule looks at (eq1 = 0: the low byte, what `.byte()` keeps;
// eq1 = 1: 1 where the element equals 1, else 0).  Non-zero byte == foreground in both cases.
PVV_HOST_LOCAL void pvv_narrow_i64(const int64_t *src, uint8_t *dst, size_t n, int eq1);
PVV_HOST_LOCAL void pvv_narrow_i32(const int32_t *src, uint8_t *dst, size_t n, int eq1);
PVV_HOST_LOCAL void pvv_narrow_f32(const float *src, uint8_t *dst, size_t n, int eq1);

// Writes the cache lines of [p, p + n) back to memory (clwb; a no-op where the CPU has none).
PVV_HOST_LOCAL void pvv_writeback(const uint8_t *p, size_t n);

// Foreground (non-zero byte) statistics of `rows` rows of `w` bytes: returns the number of foreground pixels and adds the
// summed row extents (last - first + 1 of every row with foreground) to *span_px.
PVV_HOST_LOCAL long long pvv_fg_extent(const uint8_t *bytes, int w, int rows, long long *span_px);
