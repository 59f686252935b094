"""Small end-to-end case for the checked build of the library (bounds asserts on pool / chunk-table / queue / counts indices,
-DPVV_CHECKED: see PVV_CHECK in pvnet_b200/csrc/pvv_common.cuh).  compute-sanitizer is not available on this GPU pool; the
checked build is what replaces it:
    tools/build_variant.sh checked -DPVV_CHECKED                  (CPU container)
    PVV_LIB_PATH=tools/_build/checked.so python tools/sanitize_case.py      (GPU box)
A failed check prints its site and traps (the call returns PVV_ERR_CUDA).  Covers every kernel: v3 (filter + exact kernels,
subsampling, skipped image, every K3 shape), v5, hypotheses, covariance, motion, primitives, host-buffer path."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pvnet_b200 import ransac_voting_gpu as pv, synth

dev = "cuda:0"
mask, vertex, _ = synth.make_batch(3, vn=3, frac=0.12, sigma=0.05, seed=5, h=96, w=128)
mask[1] = 0
tm, tv = torch.from_numpy(mask).to(dev), synth.as_net_output(vertex, device=dev)
for var in (0, 1, 2, 7, 8, 9, 10, 11, 12, 13):
    o1 = pv.ransac_voting_layer_v3(tm, tv, 200, 0.999, seed=1, min_num=10, max_num=800, k3_variant=var)
    o2 = pv.ransac_voting_layer_v3(tm, tv, 200, 0.999, seed=1, min_num=10, max_num=800, force_exact=True)
    assert torch.equal(o1, o2), var
for hn in (64, 200, 520):
    out, dbg = pv.ransac_voting_layer_v3(tm, tv, hn, 0.999, seed=1, min_num=10, max_num=800, return_debug=True)
    out2, dbg2 = pv.ransac_voting_layer_v3(tm, tv, hn, 0.999, seed=1, min_num=10, max_num=800, return_debug=True, force_exact=True)
    assert torch.equal(dbg["counts"], dbg2["counts"]) and torch.equal(out, out2)
k5, c5 = pv.ransac_voting_layer_v5(tm, tv, 64, seed=2)
hyp, cnt = pv.generate_hypothesis(tm, tv, 64, seed=3)
_, cov = pv.estimate_voting_distribution_with_mean(tm, tv, out, 32, 128, seed=4, min_num=10, max_num=800)
mv = pv.ransac_motion_voting(tm, tv)
hk = pv.ransac_voting_layer_v3(torch.from_numpy(mask).pin_memory(), synth.as_net_output(vertex, pin=True), 64, 0.99, seed=1, min_num=10)
torch.cuda.synchronize()
assert torch.isfinite(cov).all() and torch.isfinite(mv).all() and torch.isfinite(hk).all()
print("checked case ok", float(out.abs().sum()), float(c5.mean()))
