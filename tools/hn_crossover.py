"""Which K3 kernel wins at which hypothesis count (GPU box): configs[1]-sized batch, FP32-pipe shapes (variants 0 / 1 / 2)
against the tensor-pipe kernel (7).  Feeds the policy in pvv_launch_count."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pvnet_b200 import ransac_voting_gpu as pv, synth

dev = "cuda:0"
mask, vertex, _ = synth.make_batch(32, vn=9, frac=0.03, sigma=0.05, seed=1234)
tm, tv = torch.from_numpy(mask).to(dev), synth.as_net_output(vertex, device=dev)
for thresh in (0.999, 0.99):
    for hn in (128, 192, 256, 320, 384, 512):
        row = []
        for var in (0, 1, 2, 7):
            st = []
            for i in range(10):
                pv.ransac_voting_layer_v3(tm, tv, hn, thresh, seed=0, stage_ms=st, k3_variant=var)
            row.append((var, float(np.median([s["count"] + s["hypotheses"] for s in st[2:]]))))
        best = min(row, key=lambda r: r[1])
        print(f"thresh {thresh} hn {hn:4d}: " + "  ".join(f"v{v} {t * 1e3:6.1f} us" for v, t in row) + f"   best v{best[0]}", flush=True)
