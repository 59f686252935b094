"""Development driver for the second-generation tensor-pipe kernel (GPU box): correctness of the forced variants against
the exact-only kernel on ragged / adversarial cases, then timing at configs[1] and a few other shapes.
usage: python tools/dev_tc2.py [variants...]   (default 7 10)"""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pvnet_b200 import ransac_voting_gpu as pv, synth

dev = "cuda:0"
variants = [int(v) for v in sys.argv[1:]] or [7, 10]
rng = np.random.default_rng(777)
cases = []
for case in range(14):
    h, w = int(rng.integers(24, 300)), int(rng.integers(24, 400))
    b, vn = int(rng.integers(1, 6)), int(rng.integers(1, 12))
    hn = int(rng.choice([7, 128, 130, 300, 512, 700, 1100]))
    thresh = float(rng.choice([0.5, 0.9, 0.99, 0.999, 0.9999, 0.99999]))
    masks = np.zeros((b, h, w), np.int64)
    verts = np.zeros((b, h, w, vn, 2), np.float32)
    for i in range(b):
        r = np.random.default_rng(int(rng.integers(1 << 30)))
        masks[i] = synth.blob_mask(r, float(rng.uniform(0.01, 0.5)), h, w)
        kp = np.stack([r.uniform(-3 * w, 4 * w, vn), r.uniform(-3 * h, 4 * h, vn)], 1) if rng.random() < 0.3 else \
            np.stack([r.uniform(0, w, vn), r.uniform(0, h, vn)], 1)
        if rng.random() < 0.3:
            kp = np.round(kp)
        verts[i] = synth.unit_field(masks[i], kp, float(rng.choice([0.0, 0.02, 0.2])), r)
        if rng.random() < 0.3:
            verts[i] *= r.choice([0.0, 1e-7, 1e-3, 1.0, 1e3, 1e18], size=(h, w, vn, 1)).astype(np.float32)
    if case % 5 == 4:
        masks[0] = 0
    cases.append((torch.from_numpy(masks).to(dev), torch.from_numpy(verts).to(dev), hn, thresh, case))
ok = True
for m, v, hn, th, c in cases:
    want = pv.ransac_voting_layer_v3(m, v, hn, th, seed=c, min_num=5, max_num=10 ** 6, return_debug=True, force_exact=True)[1]["counts"].clone()
    for var in variants:
        got = pv.ransac_voting_layer_v3(m, v, hn, th, seed=c, min_num=5, max_num=10 ** 6, return_debug=True, k3_variant=var)[1]
        torch.cuda.synchronize()
        same = torch.equal(got["counts"], want)
        ok &= same
        print(f"case {c} b{m.shape[0]} {m.shape[1]}x{m.shape[2]} vn{v.shape[3]} hn{hn} t{th} var{var}: same={same} amb={int(got['amb_count'])}"
              + ("" if same else f"  DIFF {int((got['counts'] != want).sum())}"), flush=True)
print("CORRECT" if ok else "MISMATCH", flush=True)

mask, vertex, _ = synth.make_batch(32, vn=9, frac=0.03, sigma=0.05, seed=1234)
tm, tv = torch.from_numpy(mask).to(dev), synth.as_net_output(vertex, device=dev)
tests = float((mask != 0).sum()) * 9
for hn, thresh in ((512, 0.999), (512, 0.99), (128, 0.99), (256, 0.999), (2048, 0.999)):
    base = pv.ransac_voting_layer_v3(tm, tv, hn, thresh, seed=0, return_debug=True, force_exact=True)[1]["counts"].clone()
    for var in variants:
        st = []
        for i in range(14):
            out, dbg = pv.ransac_voting_layer_v3(tm, tv, hn, thresh, seed=0, stage_ms=st, return_debug=True, k3_variant=var)
        ms = {k: float(np.median([s[k] for s in st[3:]])) for k in st[0]}
        print(f"hn {hn} thresh {thresh} variant {var}: count {ms['count']*1e3:.1f} us hyp {ms['hypotheses']*1e3:.1f} total {ms['total']*1e3:.1f} us  "
              f"{tests*hn/ms['count']/1e9:.3f} T tests/s  amb {int(dbg['amb_count'])/(tests*hn):.2e}  same_counts={torch.equal(dbg['counts'], base)}", flush=True)
