"""Same-box A/B of two builds of libpvvote.so on the headline workload (GPU box; boxes differ by ~2 %, so compare builds
only within one call).  usage: python tools/ab_lib.py libA.so libB.so [thresh]   -- each runs in its own process, ABAB."""
import os, subprocess, sys
if sys.argv[1] == "--child":
    import numpy as np, torch
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from pvnet_b200 import _lib
    _lib.LIB_PATH = os.path.abspath(sys.argv[2])
    from pvnet_b200 import ransac_voting_gpu as pv, synth
    th = float(sys.argv[3])
    mask, vertex, _ = synth.make_batch(32, vn=9, frac=0.03, sigma=0.05, seed=1234)
    tm, tv = torch.from_numpy(mask).to("cuda:0"), synth.as_net_output(vertex, device="cuda:0")
    st = []
    for i in range(40):
        pv.ransac_voting_layer_v3(tm, tv, 512, th, seed=0, stage_ms=st)
    print(" ".join(f"{k}={float(np.median([s[k] for s in st[8:]])) * 1e3:.1f}" for k in st[0]))
    sys.exit(0)
a, b = sys.argv[1], sys.argv[2]
th = sys.argv[3] if len(sys.argv) > 3 else "0.999"
for rep in range(3):
    for lib in (a, b):
        r = subprocess.run([sys.executable, __file__, "--child", lib, th], capture_output=True, text=True)
        print(f"{os.path.basename(lib):28s} {r.stdout.strip() or r.stderr[-300:]}", flush=True)
