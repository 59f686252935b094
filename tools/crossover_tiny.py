"""K3 kernel choice at <= 128 hypotheses on small jobs (GPU box): the reference's shipped setting (128 hypotheses, 0.99, max_num = 100)
and small batches with full pools; K2 + K3 stage time, FP32-pipe one-warp shape (variant 2) against the tensor-pipe kernel (7)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pvnet_b200 import ransac_voting_gpu as pv, synth
dev = "cuda:0"
mask, vertex, _ = synth.make_batch(32, vn=9, frac=0.03, sigma=0.05, seed=1234)
for b, hn, max_num in ((1, 128, 100), (8, 128, 100), (32, 128, 100), (32, 64, 100), (1, 128, 30000), (2, 128, 30000), (4, 128, 30000), (1, 64, 30000), (4, 64, 30000), (32, 64, 30000)):
    tm, tv = torch.from_numpy(mask[:b]).to(dev), synth.as_net_output(vertex[:b], device=dev)
    row = []
    for var in (2, 7):
        st = []
        for i in range(14):
            pv.ransac_voting_layer_v3(tm, tv, hn, 0.99, seed=0, min_num=5, max_num=max_num, stage_ms=st, k3_variant=var)
        row.append((float(np.median([s["count"] + s["hypotheses"] for s in st[3:]])) * 1e3, float(np.median([s["total"] for s in st[3:]])) * 1e3))
    print(f"b {b:2d} hn {hn:3d} max_num {max_num:5d}: FP32-pipe K2+K3 {row[0][0]:6.1f} us (call {row[0][1]:6.1f})   tensor-pipe {row[1][0]:6.1f} us (call {row[1][1]:6.1f})", flush=True)
# what the library picks by itself
for b, hn, max_num in ((1, 128, 100), (32, 128, 100), (1, 128, 30000), (4, 128, 30000), (8, 128, 30000), (16, 128, 30000), (32, 128, 30000)):
    tm, tv = torch.from_numpy(mask[:b]).to(dev), synth.as_net_output(vertex[:b], device=dev)
    st = []
    for i in range(14):
        pv.ransac_voting_layer_v3(tm, tv, hn, 0.99, seed=0, min_num=5, max_num=max_num, stage_ms=st)
    print(f"default choice: b {b:2d} hn {hn} max_num {max_num}: K2+K3 {float(np.median([s['count'] + s['hypotheses'] for s in st[3:]])) * 1e3:6.1f} us", flush=True)
