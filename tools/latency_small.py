"""Latency of one call on small problems (GPU box): configs[0] (the demo's cat mask, b = 1) and b = 1..8 at configs[1]'s
per-image size.  CUDA events around single calls and around CUDA-graph replays of the same call."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pvnet_b200 import ransac_voting_gpu as pv, synth

dev = "cuda:0"


def timeit(fn, n=200):
    for _ in range(20):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


def graphed(fn):
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        fn()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            fn()
    torch.cuda.current_stream().wait_stream(s)
    return g.replay


mask, kp = synth.cat_fixture()
vertex = synth.unit_field(mask.astype(np.int64), kp, 0.02, np.random.default_rng(0))[None]
tm, tv = torch.from_numpy(mask[None].astype(np.int64)).to(dev), synth.as_net_output(vertex, device=dev)
for hn, th in ((128, 0.99), (512, 0.99)):
    f = lambda: pv.ransac_voting_layer_v3(tm, tv, hn, th, seed=1)
    print(f"configs[0] cat b=1 hn={hn}: {timeit(f):7.1f} us per call, {timeit(graphed(f)):7.1f} us per graph replay", flush=True)
for b in (1, 2, 4, 8):
    m, v, _ = synth.make_batch(b, vn=9, frac=0.03, sigma=0.05, seed=9)
    tmb, tvb = torch.from_numpy(m).to(dev), synth.as_net_output(v, device=dev)
    f = lambda: pv.ransac_voting_layer_v3(tmb, tvb, 512, 0.999, seed=1)
    print(f"configs[1]-sized images b={b} hn=512: {timeit(f):7.1f} us per call, {timeit(graphed(f)):7.1f} us per graph replay", flush=True)
