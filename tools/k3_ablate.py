"""Ablation timing of K3's phases on the headline workload (GPU box).  Builds a second library with -DPVV_K3_ABLATE
(tools/_build/libpvvote_ablate.so, here, before gpurun) and times k_count with phases left out; counts are WRONG then.
    python tools/k3_ablate.py build      (CPU container)
    python tools/k3_ablate.py            (GPU box)
bits: 1 = no deferred passes (push, phase A, phase B), 2 = no hot loop, 4 = no constants staging, 8 = no count atomics"""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "tools", "_build")
LIB = os.path.join(OUT, "libpvvote_ablate.so")
if sys.argv[1:] == ["build"]:
    from pvnet_b200 import build as pb
    os.makedirs(OUT, exist_ok=True)
    srcs = [os.path.join(pb.CSRC, s) for s in pb.SOURCES]
    for hv in [int(v) for v in os.environ.get('K3_BUILD_HV', '0').split(',')]:
        subprocess.run([pb.NVCC, *pb.FLAGS[:-2], "-DPVV_K3_ABLATE", f"-DK3_HOTVAR={hv}", "-shared", "-cudart", "static", *srcs,
                        "-o", LIB.replace(".so", f"_hv{hv}.so")], check=True)
    print(LIB)
    sys.exit(0)
import numpy as np, torch
from pvnet_b200 import _lib
HV = int(os.environ.get("K3_HOTVAR", "0"))
_lib.LIB_PATH = LIB.replace(".so", f"_hv{HV}.so")
from pvnet_b200 import ransac_voting_gpu as pv, synth
dev = "cuda:0"
mask, vertex, _ = synth.make_batch(32, vn=9, frac=0.03, sigma=0.05, seed=1234)
tm, tv = torch.from_numpy(mask).to(dev), synth.as_net_output(vertex, device=dev)
HN = int(os.environ.get("K3_HN", "512"))
VAR = os.environ.get("K3_VAR")
for thresh in (0.999,):
    for abl in [int(v) for v in (sys.argv[1:] or [0, 1, 4, 5, 8, 13, 2, 3, 15])]:
        os.environ["PVV_K3_ABLATE"] = str(abl)
        st = []
        for i in range(14):
            pv.ransac_voting_layer_v3(tm, tv, HN, thresh, seed=0, stage_ms=st, k3_variant=int(VAR) if VAR else None)
        print(f"hn {HN} variant {VAR} hotvar {HV} thresh {thresh} ablate {abl:2d}: count {float(np.median([s['count'] for s in st[3:]])):.4f} ms", flush=True)
