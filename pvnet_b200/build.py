"""Build the native pieces in-tree for sm_100a.

    python -m pvnet_b200.build [--force]

  libpvvote.so      the C-ABI CUDA library (include/pvvote.h).  Explicit nvcc, no torch involvement: the library has
                    no torch/ATen dependency.
  ransac_voting.so  the PyTorch C++ extension that replaces the reference's src/ransac_voting.cpp + .cu + setup.py:
                    a thin shim (csrc/ransac_voting_shim.cpp, plain g++) over libpvvote.so.
"""
from __future__ import annotations

import concurrent.futures as cf
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(CSRC, "build")
LIB = os.path.join(HERE, "libpvvote.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
SOURCES = ["pvv_compact.cu", "pvv_vote.cu", "pvv_finalize.cu", "pvv_primitives.cu", "pvv_pnp.cu", "pvv_capi.cu", "pvv_host.cu",
           "pvv_host_simd.cpp"]
HEADERS = ["pvv_common.cuh", "pvv_internal.cuh", "pvv_vote_tc.cuh", "pvv_vote_tc2.cuh", "pvv_host_simd.h",
           "../../include/pvvote.h"]
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
         "-Xcompiler", "-fPIC", "-Xptxas", "-v"]


def _stale(dst, srcs):
    if not os.path.exists(dst):
        return True
    t = os.path.getmtime(dst)
    return any(os.path.getmtime(s) > t for s in srcs)


def _compile(src):
    s = os.path.join(CSRC, src)
    o = os.path.join(OBJ, os.path.splitext(src)[0] + ".o")   # nvcc hands a .cpp to the host compiler as it is
    deps = [s] + [os.path.join(CSRC, h) for h in HEADERS]
    if not _stale(o, deps):
        return o, ""
    r = subprocess.run([NVCC, *FLAGS, "-c", s, "-o", o], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed on {src}:\n{r.stdout}\n{r.stderr}")
    return o, r.stderr


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    if force:
        for f in os.listdir(OBJ):
            os.remove(os.path.join(OBJ, f))
    with cf.ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        res = list(ex.map(_compile, SOURCES))
    objs = [o for o, _ in res]
    if verbose:
        for _, log in res:
            if log:
                print(log)
    if _stale(LIB, objs):
        r = subprocess.run([NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-cudart", "static",
                            *objs, "-o", LIB], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


SHIM = os.path.join(HERE, "ransac_voting.so")


def build_shim(force: bool = False) -> str:
    """g++ the torch extension shim against torch's headers; links libpvvote.so by rpath ($ORIGIN)."""
    src = os.path.join(CSRC, "ransac_voting_shim.cpp")
    hdr = os.path.join(HERE, "..", "include", "pvvote.h")
    if not force and not _stale(SHIM, [src, hdr, LIB]):
        return SHIM
    import sysconfig

    import torch
    from torch.utils import cpp_extension as ce

    inc = []
    for p in ce.include_paths("cuda") + [sysconfig.get_paths()["include"], "/usr/local/cuda/include"]:
        inc += ["-I", p]
    libdir = os.path.join(os.path.dirname(torch.__file__), "lib")
    abi = int(torch._C._GLIBCXX_USE_CXX11_ABI)
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-w", "-DTORCH_EXTENSION_NAME=ransac_voting",
           "-DTORCH_API_INCLUDE_EXTENSION_H", f"-D_GLIBCXX_USE_CXX11_ABI={abi}", *inc, src, "-o", SHIM,
           f"-L{HERE}", "-lpvvote", "-Wl,-rpath,$ORIGIN", f"-L{libdir}", f"-Wl,-rpath,{libdir}",
           "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch", "-ltorch_python"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"shim build failed:\n{r.stdout}\n{r.stderr}")
    return SHIM


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
    print(build_shim(force="--force" in sys.argv))
