"""Tensor-pipe K3 (variant 8 = 7 producer warps) against the FP32-pipe shapes at 128 / 256 / 512 hypotheses (GPU box)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pvnet_b200 import ransac_voting_gpu as pv, synth
dev = "cuda:0"
mask, vertex, _ = synth.make_batch(32, vn=9, frac=0.03, sigma=0.05, seed=1234)
tm, tv = torch.from_numpy(mask).to(dev), synth.as_net_output(vertex, device=dev)
px = float((mask != 0).sum())
for thresh in (0.999, 0.99):
    for hn, vs in ((128, (2, 8)), (256, (1, 8)), (512, (8,))):
        base = pv.ransac_voting_layer_v3(tm, tv, hn, thresh, seed=0, return_debug=True, force_exact=True)[1]["counts"].clone()
        for var in vs:
            st = []
            for i in range(14):
                out, dbg = pv.ransac_voting_layer_v3(tm, tv, hn, thresh, seed=0, stage_ms=st, return_debug=True, k3_variant=var)
            ms = float(np.median([s["count"] for s in st[3:]]))
            t = px * 9 * hn
            print(f"thresh {thresh} hn {hn:4d} variant {var}: count {ms * 1e3:6.1f} us  {t / ms / 1e9:.2f} T tests/s  frac {t / ms / 1e9 * 12 / 72.6:.3f}  same_counts={torch.equal(dbg['counts'], base)}", flush=True)
