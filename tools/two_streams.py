"""Throughput of back-to-back calls on 1 vs 2 streams (GPU box): does the memory-bound front (K1) of call i+1 overlap the
compute-bound K3 of call i?  Inputs rotate over 3 sets larger than L2, as in bench.py."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pvnet_b200 import ransac_voting_gpu as pv, synth
dev = "cuda:0"
sets = []
for s in range(3):
    m, v, _ = synth.make_batch(32, vn=9, frac=0.03, sigma=0.05, seed=1234 + 100 * s)
    sets.append((torch.from_numpy(m).to(dev), synth.as_net_output(v, device=dev)))


def run(nstreams, steps=200):
    streams = [torch.cuda.Stream() for _ in range(nstreams)]
    outs = []
    def loop(n):
        for i in range(n):
            with torch.cuda.stream(streams[i % nstreams]):
                outs.append(pv.ransac_voting_layer_v3(*sets[i % 3], 512, 0.999, seed=i))
            if len(outs) > 8:
                outs.pop(0)
    cur = torch.cuda.current_stream()
    for s in streams:
        s.wait_stream(cur)
    loop(20)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(cur)
    for s in streams:
        s.wait_stream(cur)
    loop(steps)
    for s in streams:
        cur.wait_stream(s)
    e1.record(cur)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


for var in ("", "6", "3"):
    pv._K3_VARIANT_ENV = var or None   # the module-level default of the k3_variant keyword
    for ns in (1, 2, 3):
        ms = run(ns)
        print(f"K3 variant {var or 'default'} streams {ns}: {ms:.4f} ms/step  {32 / ms * 1e3:.0f} images/s", flush=True)
