"""FP32-pipe K3 shapes at few hypotheses (GPU box): time, tests/s and counts against the exact-only kernel.
usage: k3_small_hn.py [hn:variant,variant,... ...]   default: 64 / 128 / 256 with the one-warp, two-warp and PSPLIT shapes"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pvnet_b200 import ransac_voting_gpu as pv, synth

dev = "cuda:0"
plan = sys.argv[1:] or ["64:2,11,12", "128:2,11,12,1", "256:1,13,0,7"]
for frac in (0.03, 0.10):
    mask, vertex, _ = synth.make_batch(32 if frac < 0.05 else 8, vn=9, frac=frac, sigma=0.05, seed=1234)
    tm, tv = torch.from_numpy(mask).to(dev), synth.as_net_output(vertex, device=dev)
    px = float((mask != 0).sum())
    for spec in plan:
        hn, vs = spec.split(":")
        hn = int(hn)
        tests = px * 9 * hn
        for thresh in (0.999, 0.99):
            base = pv.ransac_voting_layer_v3(tm, tv, hn, thresh, seed=0, return_debug=True, force_exact=True)[1]["counts"].clone()
            for var in [int(v) for v in vs.split(",")]:
                st = []
                for i in range(14):
                    out, dbg = pv.ransac_voting_layer_v3(tm, tv, hn, thresh, seed=0, stage_ms=st, return_debug=True, k3_variant=var)
                ms = float(np.median([s["count"] for s in st[3:]]))
                print(f"fg {frac} hn {hn:4d} thresh {thresh} variant {var:2d}: count {ms * 1e3:7.1f} us  {tests / ms / 1e9:.3f} T tests/s"
                      f"  frac {tests / ms / 1e9 * 12 / 72.6:.3f}  same_counts={torch.equal(dbg['counts'], base)}", flush=True)
