"""SASS evidence for the counting kernels of libpvvote.so (run here, no GPU): per kernel, how often the instructions that
prove tcgen05 / TMEM / TMA -- and the ones that betray spills -- occur.   python tools/sass_grep.py > profiles/r02_k_count_tc_sass_grep.txt"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "pvnet_b200", "libpvvote.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
pat = re.compile(r"\b(UTC[A-Z]+|LDTM|STTM|UBLKCP|UTMA[A-Z]+|SYNCS|STL|LDL|LDGSTS|USETMAXREG|ELECT|NANOSLEEP|FFMA2|FMNMX3|PRMT)\b")
cur, counts, sizes = None, collections.OrderedDict(), {}
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        counts[cur] = collections.Counter()
        sizes[cur] = 0
        continue
    if cur and re.search(r"/\*[0-9a-f]{4}\*/", line):
        sizes[cur] += 1
        for op in pat.findall(line):
            counts[cur][op] += 1
print(f"# cuobjdump -sass {os.path.relpath(lib, ROOT)} (sm_100a), instruction counts per kernel; tcgen05.mma = UTC*MMA, tcgen05.ld = LDTM,")
print("# tcgen05.commit = UTCBAR, cp.async.bulk (1-D TMA) = UBLKCP, cp.async = LDGSTS, mbarrier = SYNCS, spills = STL / LDL")
for k, c in counts.items():
    if "k_count" not in k:
        continue
    name = k.replace("(anonymous namespace)::", "").replace("void ", "").split("(")[0]
    print(f"\n{name}   [{sizes[k]} SASS instructions]")
    print("   " + "  ".join(f"{op} x{n}" for op, n in sorted(c.items())))
