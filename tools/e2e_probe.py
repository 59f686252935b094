"""Where does the host-buffer call spend its time?  (GPU box)"""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pvnet_b200 import ransac_voting_gpu as pv, synth
dev = "cuda:0"
mask, vertex, _ = synth.make_batch(32, vn=9, frac=0.03, sigma=0.05, seed=1234)
hm = torch.from_numpy(mask).pin_memory()
hv = synth.as_net_output(vertex, pin=True)
dm = torch.empty_like(hm, device=dev)
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
print("mask H2D copy (78.6 MB pinned): %.3f ms" % t(lambda: dm.copy_(hm, non_blocking=True)))
full = hv._base if hv._base is not None else hv
dfull = torch.empty(full.shape, device=dev)
print("full vertex H2D copy (%.0f MB): %.3f ms" % (full.numel() * 4 / 1e6, t(lambda: dfull.copy_(full, non_blocking=True), 5)))
print("v3 host call (int64 mask): %.3f ms" % t(lambda: pv.ransac_voting_layer_v3(hm, hv, 512, 0.999, seed=1)))
hm8 = hm.to(torch.uint8).pin_memory()
print("v3 host call (uint8 mask): %.3f ms" % t(lambda: pv.ransac_voting_layer_v3(hm8, hv, 512, 0.999, seed=1)))
tm, tv = hm.to(dev), synth.as_net_output(vertex, device=dev)
print("v3 device call: %.3f ms" % t(lambda: pv.ransac_voting_layer_v3(tm, tv, 512, 0.999, seed=1)))
# device mask, pinned vertex (zero-copy gather only)
st = []
print("v3 device mask + zero-copy pinned vertex: %.3f ms" % t(lambda: pv.ransac_voting_layer_v3(tm, hv.cuda() if False else tv, 512, 0.999, seed=1)))
