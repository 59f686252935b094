"""CPU-side checks of the drop-in boundary: libpvvote.so loads, exports exactly what include/pvvote.h declares, the
ctypes mirror of the structs matches the C layout, the host-only entry points validate their arguments, and the
product refuses to run without its CUDA library or a CUDA device (no fallback).  No kernel is launched here."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "pvvote.h")


@pytest.fixture(scope="module")
def built_lib():
    from pvnet_b200 import build as pbuild
    path = pbuild.build()
    assert os.path.exists(path)
    return path


def _declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(pvv_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_all_exported(built_lib):
    from pvnet_b200 import _lib
    names = _declared_functions()
    assert len(names) >= 17
    assert sorted(_lib.SYMBOLS) == names          # the ctypes binding lists exactly what the header declares
    L = C.CDLL(built_lib)
    for n in names:
        assert hasattr(L, n), f"libpvvote.so does not export {n}"
    out = subprocess.run(["nm", "-D", "--defined-only", built_lib], capture_output=True, text=True).stdout
    exported = set(re.findall(r"\sT\s+(pvv_[a-z0-9_]+)$", out, flags=re.M))
    assert exported == set(names), exported ^ set(names)   # and nothing undeclared leaks through the C ABI


def test_library_has_no_torch_dependency(built_lib):
    out = subprocess.run(["ldd", built_lib], capture_output=True, text=True).stdout
    assert "torch" not in out and "c10" not in out and "python" not in out


def test_ctypes_structs_match_c_layout(tmp_path, built_lib):
    """Compile a tiny C program against include/pvvote.h and compare sizeof / offsetof with the ctypes mirror."""
    from pvnet_b200 import _lib
    fields = {"pvv_job": [f for f, _ in _lib.Job._fields_], "pvv_debug": [f for f, _ in _lib.Debug._fields_]}
    lines = ["#include <stdio.h>", "#include <stddef.h>", f'#include "{HEADER}"', "int main(void){"]
    for s, fs in fields.items():
        lines.append(f'printf("{s} %zu\\n", sizeof({s}));')
        for f in fs:
            lines.append(f'printf("{s}.{f} %zu\\n", offsetof({s}, {f}));')
    lines.append("return 0;}")
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-std=c11", str(src), "-o", str(exe)], check=True)
    got = dict(l.split() for l in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.splitlines())
    for s, cls in (("pvv_job", _lib.Job), ("pvv_debug", _lib.Debug)):
        assert int(got[s]) == C.sizeof(cls)
        for f, _ in cls._fields_:
            assert int(got[f"{s}.{f}"]) == getattr(cls, f).offset, (s, f)


def _job(b=2, h=480, w=640, vn=9, hn=512, max_num=30000):
    from pvnet_b200 import _lib
    j = _lib.Job()
    j.b, j.h, j.w, j.vn, j.hn_total = b, h, w, vn, hn
    j.min_num, j.max_num, j.inlier_thresh = 100, max_num, 0.99
    return j


def test_workspace_sizing_is_host_only_and_monotone(built_lib):
    from pvnet_b200 import _lib
    L = _lib.lib()
    j = _job()
    n1 = L.pvv_workspace_bytes(C.byref(j))
    cap = L.pvv_default_pool_capacity(C.byref(j))
    assert cap % 64 == 0 and 2 * 30000 < cap < 2 * 480 * 640          # max_num bound + Bernoulli slack, per image
    assert n1 > cap * (4 + 9 * 8)                                       # pool: xy + vn float2 per pixel
    j2 = _job(b=4)
    assert L.pvv_workspace_bytes(C.byref(j2)) > n1
    j3 = _job(max_num=10 ** 9)                                         # no subsampling bound: whole images
    # no bound below the image: whole images, plus the alignment slack (every image starts on a 16-pixel boundary)
    assert 2 * 480 * 640 <= L.pvv_default_pool_capacity(C.byref(j3)) <= 2 * 480 * 640 + 16 * 3 + 64
    j4 = _job()
    j4.pool_capacity = 1000
    assert 1000 + 16 * 2 <= L.pvv_default_pool_capacity(C.byref(j4)) <= 1000 + 16 * 3 + 64
    bad = _job(b=0)
    assert L.pvv_workspace_bytes(C.byref(bad)) == 0
    assert L.pvv_workspace_bytes(None) == 0
    assert b"sm_100a" in L.pvv_version()


def test_argument_validation_returns_status_codes(built_lib):
    """Bad arguments come back as pvv_status values: nothing throws, exits or touches the device."""
    from pvnet_b200 import _lib
    L = _lib.lib()
    out = (C.c_float * 64)()
    assert L.pvv_v3(None, 0.99, 100, out, None) == -1
    j = _job()
    assert L.pvv_v3(C.byref(j), 0.99, 100, out, None) == -1             # null mask / vertex / workspace
    assert L.pvv_v3(C.byref(j), 0.99, 100, None, None) == -1
    assert L.pvv_cov(C.byref(j), None, out, None) == -1
    buf = (C.c_char * 4096)()
    j.mask, j.vertex, j.workspace, j.workspace_bytes = C.addressof(buf), C.addressof(buf), C.addressof(buf), 4096
    j.mask_dtype = 7
    assert L.pvv_v3(C.byref(j), 0.99, 100, out, None) == -1             # unknown mask dtype
    j.mask_dtype, j.vn = 0, 65
    assert L.pvv_v3(C.byref(j), 0.99, 100, out, None) == -4             # vn above the compiled limit
    j.vn = 9
    assert L.pvv_v3(C.byref(j), 0.99, 100, out, None) == -2             # workspace too small
    j.inlier_thresh = float("nan")
    assert L.pvv_v3(C.byref(j), 0.99, 100, out, None) == -1
    assert L.pvv_generate_hypothesis(None, None, None, 1, 1, 1, None, None) == -1
    assert L.pvv_voting_for_hypothesis(None, None, None, None, 1, 1, 1, 0.9, None) == -1
    assert L.pvv_vote_counts(None, None, None, 1, 1, 1, 0.9, None, None) == -1
    assert L.pvv_make_idxs(0, 0, 1, 1, 1, None, None, None) == -1
    assert L.pvv_make_selection(0, 0, 1, 1, 1, None, None) == -1
    with pytest.raises(RuntimeError, match="PVV_ERR_INVALID"):
        _lib.check(-1, "x")


def test_missing_library_fails_loudly(monkeypatch, built_lib):
    from pvnet_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", os.path.join(ROOT, "pvnet_b200", "does_not_exist.so"))
    with pytest.raises(RuntimeError, match="no CPU or PyTorch fallback"):
        _lib.lib()


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_no_cuda_device_means_error_not_fallback(built_lib):
    from pvnet_b200 import ransac_voting_gpu as pv
    mask = torch.ones((1, 8, 8), dtype=torch.int64)
    vertex = torch.zeros((1, 8, 8, 2, 2))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        pv.ransac_voting_layer_v3(mask, vertex, 16)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        pv.estimate_voting_distribution_with_mean(mask, vertex, torch.zeros((1, 2, 2)))
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        pv.ransac_voting.generate_hypothesis(torch.zeros(4, 2, 2), torch.zeros(4, 2), torch.zeros(3, 2, 2, dtype=torch.int32))


def test_product_never_imports_oracle():
    """The product path must not route through the checker: nothing under pvnet_b200/ mentions oracle."""
    pkg = os.path.join(ROOT, "pvnet_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", txt, flags=re.M), f
                assert "liborc" not in txt and "pvv_oracle" not in txt, f


def test_signatures_match_reference_api():
    """Same positional names and defaults as ransac_voting_gpu.py:520-521 and :333-334."""
    import inspect
    from pvnet_b200 import ransac_voting_gpu as pv
    s = inspect.signature(pv.ransac_voting_layer_v3)
    pos = [(n, p.default) for n, p in s.parameters.items() if p.kind == p.POSITIONAL_OR_KEYWORD]
    assert pos == [("mask", inspect._empty), ("vertex", inspect._empty), ("round_hyp_num", inspect._empty),
                   ("inlier_thresh", 0.99), ("confidence", 0.99), ("max_iter", 100), ("min_num", 100),
                   ("max_num", 30000)]
    s = inspect.signature(pv.estimate_voting_distribution_with_mean)
    pos = [(n, p.default) for n, p in s.parameters.items() if p.kind == p.POSITIONAL_OR_KEYWORD]
    assert pos == [("mask", inspect._empty), ("vertex", inspect._empty), ("mean", inspect._empty),
                   ("round_hyp_num", 256), ("min_hyp_num", 4096), ("topk", 128), ("inlier_thresh", 0.99),
                   ("min_num", 20), ("max_num", 30000), ("output_hyp", False)]


def test_torch_extension_shim_builds_and_mirrors_the_reference_pybind_module(built_lib):
    """pvnet_b200/ransac_voting.so replaces the reference's pybind module (src/ransac_voting.cpp:102-107): same module
    name, same two primitive names, same CHECK_CUDA / CHECK_CONTIGUOUS errors."""
    from pvnet_b200 import build as pbuild
    from pvnet_b200 import ransac_voting_gpu as pv
    assert os.path.exists(pbuild.build_shim())
    ext = pv.torch_extension()
    assert ext.__name__ == "ransac_voting"
    for name in ("generate_hypothesis", "voting_for_hypothesis", "ransac_voting_layer_v3",
                 "estimate_voting_distribution_with_mean"):
        assert callable(getattr(ext, name))
    d, c, i = torch.zeros(4, 2, 2), torch.zeros(4, 2), torch.zeros(3, 2, 2, dtype=torch.int32)
    with pytest.raises(RuntimeError, match="must be a CUDA tensor"):
        ext.generate_hypothesis(d, c, i)
    with pytest.raises(RuntimeError, match="must be a CUDA tensor"):
        ext.voting_for_hypothesis(d, c, torch.zeros(3, 2, 2), torch.zeros(3, 2, 4, dtype=torch.uint8), 0.99)
    with pytest.raises(RuntimeError, match="must be a CUDA tensor"):
        ext.ransac_voting_layer_v3(torch.ones(1, 8, 8, dtype=torch.int64), torch.zeros(1, 8, 8, 2, 2), 16)


def test_integration_stub_reexports_the_path():
    import importlib.util
    spec = importlib.util.spec_from_file_location("ransac_voting_gpu_stub", os.path.join(ROOT, "integration", "ransac_voting_gpu.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    from pvnet_b200 import ransac_voting_gpu as pv
    assert m.ransac_voting_layer_v3 is pv.ransac_voting_layer_v3
    assert m.estimate_voting_distribution_with_mean is pv.estimate_voting_distribution_with_mean
    assert hasattr(m.ransac_voting, "generate_hypothesis") and hasattr(m.ransac_voting, "voting_for_hypothesis")


def test_logits_job_and_eval_wrapper_surface(built_lib):
    """SURVEY 8(f) N3 boundary: seg_pred [b,C,h,w] is marshalled as PVV_MASK_LOGITS_F32 with the (b,h,w) strides and the
    class stride split out; EvalWrapper.forward has the reference's signature (tools/train_linemod.py:117)."""
    import inspect
    from pvnet_b200 import _lib
    from pvnet_b200 import ransac_voting_gpu as pv
    from pvnet_b200.eval_wrapper import EvalWrapper
    seg = torch.zeros((2, 3, 8, 16)).permute(0, 2, 3, 1).contiguous().permute(0, 3, 1, 2)   # NHWC strides
    vertex = torch.zeros((2, 8, 16, 4, 2))
    j = pv._make_job(pv._prep_mask(seg), vertex, 16, 0.99, 10, 100, _lib.FG_BYTE, None, None, 0, 0, 0)
    assert j.mask_dtype == _lib.MASK_LOGITS_F32 == 4 and j.seg_classes == 3 and j.seg_class_stride == 1
    assert list(j.mask_strides) == [8 * 16 * 3, 16 * 3, 3]
    L = _lib.lib()
    assert L.pvv_workspace_bytes(C.byref(j)) > 0
    buf = (C.c_char * 4096)()
    j.workspace, j.workspace_bytes = C.addressof(buf), 4096
    out = (C.c_float * 64)()
    j.seg_classes = 0
    assert L.pvv_v3(C.byref(j), 0.99, 100, out, None) == -1            # logits need a class count
    with pytest.raises(ValueError, match="seg_pred"):
        pv._make_job(torch.zeros((2, 3, 8, 15)), vertex, 16, 0.99, 10, 100, _lib.FG_BYTE, None, None, 0, 0, 0)
    sig = inspect.signature(EvalWrapper.forward)
    assert [(n, p.default) for n, p in list(sig.parameters.items())[1:]] == [
        ("seg_pred", inspect._empty), ("vertex_pred", inspect._empty), ("use_argmax", True), ("use_uncertainty", False)]


def test_host_worker_policy_of_the_python_mirror():
    """pvnet_b200/_lib.py divides the host's hardware threads between the ranks of one box (LOCAL_WORLD_SIZE) and lets
    PVV_HOST_WORKERS override; pvv_set_host_workers returns the previous setting (no GPU needed)."""
    import subprocess
    import sys
    code = ("import os; from pvnet_b200 import _lib; L = _lib.lib(); print(L.pvv_set_host_workers(-1), os.cpu_count())")

    def run(env):
        e = {k: v for k, v in os.environ.items() if k not in ("LOCAL_WORLD_SIZE", "PVV_HOST_WORKERS")}
        e.update(env)
        out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=ROOT, env=e, check=True).stdout.split()
        return int(out[0]), int(out[1])

    prev, ncpu = run({})
    assert prev == -1                                    # one rank: the library's own default
    assert run({"PVV_HOST_WORKERS": "5"})[0] == 5
    assert run({"PVV_HOST_WORKERS": "0", "LOCAL_WORLD_SIZE": "2"})[0] == 0
    for ranks in (2, 4, 8, 64):
        share = ncpu // ranks
        assert run({"LOCAL_WORLD_SIZE": str(ranks)})[0] == (min(share, 16) if share >= 3 else 0)
