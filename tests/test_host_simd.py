"""Host-side helpers of the host-buffer entry points (pvnet_b200/csrc/pvv_host_simd.cpp): narrowing of wide masks and
the foreground window, built on their own and checked against numpy.  No GPU."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def simd(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("simd") / "host_simd.so")
    subprocess.run(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-DPVV_HOST_SIMD_TEST",
                    os.path.join(ROOT, "pvnet_b200", "csrc", "pvv_host_simd.cpp"), "-o", out], check=True)
    L = C.CDLL(out)
    for name in ("pvv_narrow_i64", "pvv_narrow_i32", "pvv_narrow_f32"):
        getattr(L, name).argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
        getattr(L, name).restype = None
    L.pvv_fg_extent.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_longlong)]
    L.pvv_fg_extent.restype = C.c_longlong
    return L


@pytest.mark.parametrize("n", [0, 1, 7, 8, 63, 64, 65, 1000, 4097])
@pytest.mark.parametrize("eq1", [0, 1])
def test_narrowing_matches_numpy(simd, n, eq1):
    rng = np.random.default_rng(n + eq1)
    vals = np.array([0, 1, 2, 255, 256, 257, -1, 1 << 40, (1 << 40) + 1, -(1 << 33)], dtype=np.int64)
    for dt, fn in ((np.int64, simd.pvv_narrow_i64), (np.int32, simd.pvv_narrow_i32), (np.float32, simd.pvv_narrow_f32)):
        if dt == np.float32:
            src = rng.choice(np.array([0.0, 1.0, 2.0, 0.5, 1.5, 200.0], dtype=np.float32), size=n)
            want = (src == 1).astype(np.uint8) if eq1 else src.astype(np.int64).astype(np.uint8)
        else:
            src = rng.choice(vals, size=n).astype(dt)   # int32: wraps like a C cast, still a valid test vector
            want = (src == 1).astype(np.uint8) if eq1 else (src.astype(np.int64) & 0xff).astype(np.uint8)
        src = np.ascontiguousarray(src)
        dst = np.full(n + 16, 0xEE, dtype=np.uint8)
        fn(src.ctypes.data, dst.ctypes.data, n, eq1)
        assert np.array_equal(dst[:n], want)
        assert (dst[n:] == 0xEE).all()          # nothing written past the end


@pytest.mark.parametrize("w", [8, 13, 64, 640])
def test_foreground_extent(simd, w):
    rng = np.random.default_rng(w)
    for trial in range(20):
        rows = int(rng.integers(1, 40))
        m = np.zeros((rows, w), dtype=np.uint8)
        k = int(rng.integers(0, 6))
        for _ in range(k):
            m[rng.integers(0, rows), rng.integers(0, w)] = rng.integers(1, 256)
        if trial % 5 == 4:
            m[rng.integers(0, rows):, : rng.integers(1, w + 1)] = 1
        span = C.c_longlong(7)
        cnt = simd.pvv_fg_extent(m.ctypes.data, w, rows, C.byref(span))
        want = sum(int(np.nonzero(r)[0].max() - np.nonzero(r)[0].min() + 1) for r in m if r.any())
        assert cnt == int((m != 0).sum()) and span.value == 7 + want
