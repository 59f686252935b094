"""Timeline of the host-buffer call (GPU box; library built with -DPVV_HOST_TRACE: tools/build_variant.sh trace -DPVV_HOST_TRACE).
usage: PVV_LIB_PATH=tools/_build/trace.so python tools/host_trace.py"""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pvnet_b200 import ransac_voting_gpu as pv, synth
mask, vertex, _ = synth.make_batch(32, vn=9, frac=0.03, sigma=0.05, seed=1234)
hm = torch.from_numpy(mask).pin_memory()
if os.environ.get("TRACE_U8"):
    hm = hm.to(torch.uint8).pin_memory()
hv = synth.as_net_output(vertex, pin=True)
ys, xs = np.nonzero(mask[0])
print("image 0: fg", len(ys), "bbox", ys.min(), ys.max(), xs.min(), xs.max(), "bbox area / fg = %.2f" % ((ys.max() - ys.min() + 1) * (xs.max() - xs.min() + 1) / len(ys)), flush=True)
for i in range(6):
    t0 = time.perf_counter()
    pv.ransac_voting_layer_v3(hm, hv, int(os.environ.get('TRACE_HN', '512')), 0.999, seed=1)
    torch.cuda.synchronize()
    print("call %d: %.3f ms" % (i, (time.perf_counter() - t0) * 1e3), file=sys.stderr, flush=True)
