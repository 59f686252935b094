"""K3 kernel choice between 128 and 256 hypotheses by job size (GPU box): FP32-pipe shapes against the tensor-pipe kernel
for b = 1 .. 32 images of configs[1]'s per-image size; stage time of K2 + K3 (the tensor-pipe kernel's pre-pass rides in K2)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pvnet_b200 import ransac_voting_gpu as pv, synth
dev = "cuda:0"
mask, vertex, _ = synth.make_batch(32, vn=9, frac=0.03, sigma=0.05, seed=1234)
for b in (1, 2, 4, 8, 16, 32):
    tm, tv = torch.from_numpy(mask[:b]).to(dev), synth.as_net_output(vertex[:b], device=dev)
    for hn, fp32_var in ((128, 2), (160, 1), (192, 1), (256, 1)):
        row = []
        for var in (fp32_var, 7):
            st = []
            for i in range(12):
                pv.ransac_voting_layer_v3(tm, tv, hn, 0.999, seed=0, stage_ms=st, k3_variant=var)
            row.append(float(np.median([s["count"] + s["hypotheses"] for s in st[3:]])) * 1e3)
        print(f"b {b:2d} hn {hn:3d}: FP32-pipe (variant {fp32_var}) {row[0]:6.1f} us   tensor-pipe {row[1]:6.1f} us   {'tensor' if row[1] < row[0] else 'fp32'}", flush=True)
