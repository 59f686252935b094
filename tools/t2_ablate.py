"""Ablation timing of k_count_tc2 (GPU box).  Build first, here:  tools/build_variant.sh abl -DPVV_K3_ABLATE
then under gpurun:  PVV_LIB_PATH=tools/_build/abl.so python tools/t2_ablate.py [bits...]      (counts are WRONG when ablated)
bits: 1 no vote, 2 no TMEM loads, 4 builders publish without building B, 8 no MMA (plain arrive), 16 no queue pushes"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pvnet_b200 import ransac_voting_gpu as pv, synth
dev = "cuda:0"
mask, vertex, _ = synth.make_batch(32, vn=9, frac=0.03, sigma=0.05, seed=1234)
tm, tv = torch.from_numpy(mask).to(dev), synth.as_net_output(vertex, device=dev)
var = int(os.environ.get("T2_VARIANT", "10"))
for thresh in (0.999,):
    for abl in [int(v) for v in (sys.argv[1:] or [0, 16, 17, 19, 20, 23, 24, 27, 31])]:
        os.environ["PVV_K3_ABLATE"] = str(abl << 16)
        st = []
        for i in range(12):
            pv.ransac_voting_layer_v3(tm, tv, 512, thresh, seed=0, stage_ms=st, k3_variant=var)
        print(f"thresh {thresh} ablate {abl:2d}: count {float(np.median([s['count'] for s in st[3:]])) * 1e3:.1f} us", flush=True)
