"""Same-box comparison of several builds of libpvvote.so on the headline workload (GPU box): each build in its own process,
round-robin, 2 rounds.  usage: python tools/ab_many.py libA.so libB.so ... [--thresh 0.999]"""
import os, subprocess, sys
here = os.path.dirname(os.path.abspath(__file__))
args = sys.argv[1:]
th = "0.999"
if "--thresh" in args:
    i = args.index("--thresh"); th = args[i + 1]; del args[i:i + 2]
for rep in range(2):
    for lib in args:
        r = subprocess.run([sys.executable, os.path.join(here, "ab_lib.py"), "--child", lib, th], capture_output=True, text=True)
        print(f"{os.path.basename(lib):34s} {r.stdout.strip() or r.stderr[-300:]}", flush=True)
