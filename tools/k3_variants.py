"""Experiment driver (GPU box): time K3 variants on the headline workload and run the FP32 probes."""
import ctypes, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pvnet_b200 import ransac_voting_gpu as pv, synth, _lib

dev = "cuda:0"
mask, vertex, _ = synth.make_batch(32, vn=9, frac=0.03, sigma=0.05, seed=1234)
tm, tv = torch.from_numpy(mask).to(dev), synth.as_net_output(vertex, device=dev)
tests = float((mask != 0).sum()) * 9 * 512
L = _lib.lib()
print("probe FFMA  TF:", L.pvv_probe_tflops(0, 4096, None), " FFMA2 TF:", L.pvv_probe_tflops(1, 4096, None))
for thresh in (0.999, 0.99):
    base = (thresh, pv.ransac_voting_layer_v3(tm, tv, 512, thresh, seed=0, return_debug=True, force_exact=True)[1]["counts"].clone())
    for var in [int(v) for v in (sys.argv[1:] or ["0", "1", "2", "3", "4", "5", "6", "7"])]:
        st = []
        for i in range(12):
            out, dbg = pv.ransac_voting_layer_v3(tm, tv, 512, thresh, seed=0, stage_ms=st, return_debug=True, k3_variant=var)
        ms = float(np.median([s["count"] for s in st[2:]]))
        c = dbg["counts"].clone()
        print(f"thresh {thresh} variant {var}: count {ms:.4f} ms  {tests/ms/1e9:.3f} T tests/s  amb {int(dbg['amb_count'])/tests:.2e}"
              f"  same_counts={torch.equal(c, base[1])}")
