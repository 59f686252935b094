"""Instruction and sample shares per CUDA source line, SASS rows aggregated (inline-asm lines included), hot loop vs rest.
usage: python tools/ncu_nonhot.py report.ncu-rep first_hot_line last_hot_line [top_n]"""
import csv, io, subprocess, sys
rep, h0, h1 = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
top = int(sys.argv[4]) if len(sys.argv) > 4 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
fname = hdr = cur = None
agg = {}
for r in csv.reader(io.StringIO(out)):
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]
    elif r[0] == "Line No":
        hdr = {h: i for i, h in enumerate(r)}
    elif r[0].isdigit() and hdr:
        cur = (fname, int(r[0]))
    elif r[0] == "" and hdr and cur and len(r) > hdr["Instructions Executed"]:
        try:
            ins, smp = float(r[hdr["Instructions Executed"]] or 0), float(r[hdr["# Samples"]] or 0)
        except ValueError:
            continue
        a = agg.setdefault(cur, [0.0, 0.0])
        a[0] += ins; a[1] += smp
ti, ts = sum(a[0] for a in agg.values()), sum(a[1] for a in agg.values())
asm_lines = set()
src = open("pvnet_b200/csrc/pvv_vote.cu").read().split("\n")
for i, l in enumerate(src, 1):
    if "fma.rn.f32x2" in l or "prmt.b32" in l or "mov.b64" in l:
        asm_lines.add(i)
hot = lambda k: k[0] == "pvv_vote.cu" and (h0 <= k[1] <= h1 or k[1] in asm_lines)
nh = [(k, v) for k, v in agg.items() if not hot(k)]
print(f"total {int(ti)} warp-instr, {int(ts)} samples; non-hot: {sum(v[0] for _, v in nh) / ti:.1%} of instr, {sum(v[1] for _, v in nh) / ts:.1%} of samples")
for k, v in sorted(nh, key=lambda kv: -kv[1][0])[:top]:
    line = src[k[1] - 1].strip()[:84] if k[0] == "pvv_vote.cu" else ""
    print(f"{v[0] / ti:6.2%} ins {v[1] / ts:6.2%} smp  {k[0]}:{k[1]:<4d} {line}")
