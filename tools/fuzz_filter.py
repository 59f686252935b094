"""Randomised check of the guard-banded filter against the exact-only kernel (GPU box): counts, argmax and keypoints must
be bit-equal for every case.  Shapes, thresholds, field noise, keypoint placement (inside / far outside the image),
degenerate directions and mask dtypes are drawn at random.   python tools/fuzz_filter.py [cases] [seed]"""
import os, sys, json
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pvnet_b200 import ransac_voting_gpu as pv, synth

dev = "cuda:0"
ncase = int(sys.argv[1]) if len(sys.argv) > 1 else 150
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 7)
bad, tests_total, amb_total = [], 0.0, 0
for c in range(ncase):
    h, w = int(rng.integers(24, 481)), int(rng.integers(24, 641))
    b, vn = int(rng.integers(1, 5)), int(rng.integers(1, 18))
    hn = int(rng.choice([5, 33, 128, 200, 512, 700, 1536]))
    thresh = float(rng.choice([0.5, 0.8, 0.9, 0.99, 0.995, 0.999, 0.9995, 0.9999, 0.99999]))
    frac, sigma = float(rng.uniform(0.02, 0.5)), float(rng.choice([0.0, 0.01, 0.05, 0.3]))
    masks = np.zeros((b, h, w), np.int64)
    verts = np.zeros((b, h, w, vn, 2), np.float32)
    for i in range(b):
        r = np.random.default_rng(int(rng.integers(1 << 30)))
        masks[i] = synth.blob_mask(r, frac, h, w)
        far = rng.random() < 0.3
        kp = np.stack([r.uniform(-5 * w, 6 * w, vn), r.uniform(-5 * h, 6 * h, vn)], 1) if far else \
            np.stack([r.uniform(0, w, vn), r.uniform(0, h, vn)], 1)
        if rng.random() < 0.3:
            kp = np.round(kp)                          # keypoints exactly on pixel centres: d = 0 cases
        verts[i] = synth.unit_field(masks[i], kp, sigma, r)
        if rng.random() < 0.3:                         # unnormalised / tiny / zero / huge directions
            scale = r.choice([0.0, 1e-7, 1e-3, 1.0, 1e3, 1e18], size=(h, w, vn, 1)).astype(np.float32)
            verts[i] *= scale
    tm, tv = torch.from_numpy(masks).to(dev), synth.as_net_output(verts, device=dev)
    kw = dict(seed=int(rng.integers(1 << 30)), min_num=int(rng.choice([1, 20, 100])), max_num=int(rng.choice([300, 5000, 30000])),
              return_debug=True)
    o1, d1 = pv.ransac_voting_layer_v3(tm, tv, hn, thresh, **kw)
    o2, d2 = pv.ransac_voting_layer_v3(tm, tv, hn, thresh, force_exact=True, **kw)
    same = torch.equal(d1["counts"], d2["counts"]) and torch.equal(d1["win_idx"], d2["win_idx"]) and \
        torch.equal(o1.view(torch.int32), o2.view(torch.int32))
    tests_total += float(d1["tn"].sum()) * vn * hn
    amb_total += int(d1["amb_count"])
    if not same:
        bad.append({"case": c, "h": h, "w": w, "b": b, "vn": vn, "hn": hn, "thresh": thresh, "sigma": sigma,
                    "diff": int((d1["counts"] != d2["counts"]).sum())})
        print("MISMATCH", bad[-1], flush=True)
print(json.dumps({"cases": ncase, "mismatches": len(bad), "tests": tests_total, "in_band_tests": amb_total,
                  "in_band_fraction": amb_total / max(tests_total, 1)}))
sys.exit(1 if bad else 0)
