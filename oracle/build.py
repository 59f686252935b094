"""Build recipe for the checker (TEST INFRASTRUCTURE ONLY -- see oracle/README.md).

  build_oracle()  gcc   oracle/pvv_oracle.c                    -> oracle/liborc.so
  build_ref()     nvcc  /root/reference/lib/ransac_voting_gpu_layer/src/*.{cu,cpp}
                        (compiled where they lie, nothing copied)  -> oracle/_ref/ref_ransac_voting.so

The second target is the reference's own pybind module `ransac_voting`
(src/ransac_voting.cpp:102-107) cross-compiled for sm_100a with nvcc's default
floating-point flags, exactly as the reference's setup.py:4-15 would build it
(no -use_fast_math, -fmad on).  The single source incompatibility with torch
2.11 -- `extern THCState* state;` at src/ransac_voting.cpp:5, a type that no
longer exists -- is neutralised from the command line with -DTHCState=void, so
the sources are compiled unmodified.  Outputs go to oracle/_ref/ only
(git-ignored, travels to the GPU box with the snapshot).

Run:  python oracle/build.py [oracle|ref|all]
"""
from __future__ import annotations

import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference/lib/ransac_voting_gpu_layer/src"
REF_OUT = os.path.join(HERE, "_ref")
ORC_SO = os.path.join(HERE, "liborc.so")
REF_SO = os.path.join(REF_OUT, "ref_ransac_voting.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")


def _run(cmd):
    print("+", " ".join(cmd), flush=True)
    subprocess.run(cmd, check=True)


def _newer(dst, srcs):
    if not os.path.exists(dst):
        return False
    t = os.path.getmtime(dst)
    return all(os.path.getmtime(s) <= t for s in srcs if os.path.exists(s))


def build_oracle(force: bool = False) -> str:
    src = os.path.join(HERE, "pvv_oracle.c")
    if not force and _newer(ORC_SO, [src]):
        return ORC_SO
    _run(["gcc", "-O2", "-std=c11", "-fPIC", "-shared", "-fopenmp",
          "-ffp-contract=off", "-fno-fast-math", "-fno-math-errno",
          src, "-o", ORC_SO, "-lm"])
    return ORC_SO


def ref_sources_present() -> bool:
    return os.path.isfile(os.path.join(REF_SRC, "ransac_voting_kernel.cu"))


def build_ref(force: bool = False) -> str | None:
    """Compile the reference's CUDA voting extension for sm_100a. Returns the
    path, or None when /root/reference is absent (GPU box: prebuilt file is used)."""
    if not ref_sources_present():
        return REF_SO if os.path.exists(REF_SO) else None
    cu = os.path.join(REF_SRC, "ransac_voting_kernel.cu")
    cpp = os.path.join(REF_SRC, "ransac_voting.cpp")
    if not force and _newer(REF_SO, [cu, cpp, os.path.join(REF_SRC, "cuda_common.h")]):
        return REF_SO
    import torch
    from torch.utils import cpp_extension as ce

    os.makedirs(REF_OUT, exist_ok=True)
    inc = []
    for p in ce.include_paths("cuda") + [sysconfig.get_paths()["include"], REF_SRC]:
        inc += ["-I", p]
    abi = int(torch._C._GLIBCXX_USE_CXX11_ABI)
    common = ["-DTORCH_EXTENSION_NAME=ref_ransac_voting", "-DTORCH_API_INCLUDE_EXTENSION_H",
              f"-D_GLIBCXX_USE_CXX11_ABI={abi}", "-std=c++17"]
    o_cu = os.path.join(REF_OUT, "ransac_voting_kernel.o")
    o_cpp = os.path.join(REF_OUT, "ransac_voting.o")
    # reference setup.py passes no nvcc flags: defaults (-fmad=true, IEEE sqrt/div)
    _run([NVCC, "-gencode", "arch=compute_100a,code=sm_100a", "-w", *common, *inc,
          "-Xcompiler", "-fPIC", "-c", cu, "-o", o_cu])
    _run(["g++", "-O2", "-w", "-fPIC", "-DTHCState=void", *common, *inc, "-c", cpp, "-o", o_cpp])
    libdir = os.path.join(os.path.dirname(torch.__file__), "lib")
    _run(["g++", "-shared", o_cu, o_cpp, "-o", REF_SO,
          f"-L{libdir}", f"-Wl,-rpath,{libdir}", "-L/usr/local/cuda/lib64",
          "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch", "-ltorch_python",
          "-lcudart"])
    return REF_SO


REF_PNP_SRC = "/root/reference/lib/utils/extend_utils/src/uncertainty_pnp.cpp"
REF_PNP_INC = "/root/reference/lib/utils/extend_utils/include"
REF_PNP_SO = os.path.join(REF_OUT, "ref_pnp_functor.so")


def build_ref_pnp(force: bool = False) -> str | None:
    """The reference's uncertainty-PnP COST functor (uncertainty_pnp.cpp:7-33) compiled against the vendored header-only
    ceres/jet.h + ceres/rotation.h + Eigen, behind oracle/ref_pnp_driver.cpp.  The struct is cut out of the reference
    source at build time into oracle/_ref/pnp_functor.inc (never committed): everything from `struct
    ReprojectionErrorArray {` up to the factory comment, plus the member declarations -- the AutoDiffCostFunction factory is
    left out because <ceres/ceres.h> needs glog headers, which this image does not have.  Returns None when
    /root/reference is absent (GPU box: the prebuilt file is used)."""
    if not os.path.isfile(REF_PNP_SRC):
        return REF_PNP_SO if os.path.exists(REF_PNP_SO) else None
    drv = os.path.join(HERE, "ref_pnp_driver.cpp")
    if not force and _newer(REF_PNP_SO, [REF_PNP_SRC, drv]):
        return REF_PNP_SO
    os.makedirs(REF_OUT, exist_ok=True)
    lines = open(REF_PNP_SRC).read().split("\n")
    a = next(i for i, l in enumerate(lines) if l.startswith("struct ReprojectionErrorArray"))
    b = next(i for i, l in enumerate(lines) if "// Factory to hide" in l)
    c = next(i for i, l in enumerate(lines) if i > b and l.strip().startswith("double x2d,y2d;"))
    d = next(i for i, l in enumerate(lines) if i > c and l.strip() == "};")
    with open(os.path.join(REF_OUT, "pnp_functor.inc"), "w") as f:
        f.write("// generated by oracle/build.py from %s lines %d-%d and %d-%d; not committed\n" % (REF_PNP_SRC, a + 1, b, c + 1, d + 1))
        f.write("\n".join(lines[a:b] + lines[c:d + 1]) + "\n")
    _run(["g++", "-O2", "-w", "-std=c++14", "-fPIC", "-shared", "-I", REF_PNP_INC, "-I", os.path.join(REF_PNP_INC, "eigen3"),
          "-I", REF_OUT, drv, "-o", REF_PNP_SO])
    return REF_PNP_SO


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    if what in ("oracle", "all"):
        print(build_oracle(force=True))
    if what in ("ref", "all"):
        print(build_ref(force=True))
        print(build_ref_pnp(force=True))
