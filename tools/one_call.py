"""A few calls of v3 at configs[1] size with a forced K3 variant (for ncu).  usage: one_call.py variant [hn thresh calls]"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pvnet_b200 import ransac_voting_gpu as pv, synth
var = int(sys.argv[1]) if len(sys.argv) > 1 else 10
hn = int(sys.argv[2]) if len(sys.argv) > 2 else 512
th = float(sys.argv[3]) if len(sys.argv) > 3 else 0.999
calls = int(sys.argv[4]) if len(sys.argv) > 4 else 4
mask, vertex, _ = synth.make_batch(32, vn=9, frac=0.03, sigma=0.05, seed=1234)
tm, tv = torch.from_numpy(mask).to("cuda:0"), synth.as_net_output(vertex, device="cuda:0")
for i in range(calls):
    out = pv.ransac_voting_layer_v3(tm, tv, hn, th, seed=0, k3_variant=var if var >= 0 else None)
torch.cuda.synchronize()
print("ok", float(out.sum()))
