"""One small case through a forced K3 variant (development; use with PVV_LIB_PATH=tools/_build/dbg.so).
usage: dev_tiny.py variant [b hn thresh [h w vn frac]]"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pvnet_b200 import ransac_voting_gpu as pv, synth
a = sys.argv[1:]
var = int(a[0]) if a else 10
b, hn, th = (int(a[1]), int(a[2]), float(a[3])) if len(a) > 3 else (2, 512, 0.99)
h, w, vn, frac = (int(a[4]), int(a[5]), int(a[6]), float(a[7])) if len(a) > 7 else (480, 640, 9, 0.03)
mask, vertex, _ = synth.make_batch(b, vn=vn, frac=frac, sigma=0.05, seed=5, h=h, w=w)
tm, tv = torch.from_numpy(mask).to("cuda:0"), synth.as_net_output(vertex, device="cuda:0")
want = pv.ransac_voting_layer_v3(tm, tv, hn, th, seed=1, min_num=5, return_debug=True, force_exact=True)[1]["counts"].clone()
out, dbg = pv.ransac_voting_layer_v3(tm, tv, hn, th, seed=1, min_num=5, return_debug=True, k3_variant=var)
torch.cuda.synchronize()
print("same", torch.equal(dbg["counts"], want), "diff", int((dbg["counts"] != want).sum()), "of", want.numel(), "amb", int(dbg["amb_count"]))
