"""Per CUDA source line: stall samples and executed warp instructions of one kernel, from an ncu report (read here).
usage: python tools/ncu_lines.py report.ncu-rep [top_n]"""
import csv, subprocess, sys, io
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
fname, lines, hdr = None, [], None
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]
    elif r[0] == "Line No":
        hdr = {h: i for i, h in enumerate(r)}
    elif r[0] == "Function Name":
        func = r[1]
    elif r[0] not in ("",) and hdr and len(r) > 8 and r[0].isdigit():
        try:
            lines.append((fname, int(r[0]), r[1].strip(), float(r[hdr["# Samples"]] or 0), float(r[hdr["Instructions Executed"]] or 0)))
        except (ValueError, IndexError):
            pass
ts, ti = sum(l[3] for l in lines), sum(l[4] for l in lines)
print(f"{len(lines)} source lines with code; {int(ts)} samples, {int(ti)} warp-instr")
agg = {}
for f, n, s, smp, ins in lines:
    k = (f, n)
    a = agg.setdefault(k, [s, 0.0, 0.0])
    a[1] += smp; a[2] += ins
for (f, n), (s, smp, ins) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"{smp/ts:6.2%} smp {ins/ti:6.2%} ins  {f}:{n:<4d} {s[:90]}")
