"""Summarise an ncu report (read here, no GPU): key raw metrics + per-region stall breakdown from the source page.
usage: python tools/ncu_summary.py gpurun_out/x.ncu-rep [launch_index]"""
import collections, csv, io, subprocess, sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
r = rows[2 + (int(sys.argv[2]) if len(sys.argv) > 2 else 0)]
want = ["Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_static", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.avg", "smsp__inst_executed.sum",
        "sm__issue_active.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sass__inst_executed_register_spilling"]
for i, h in enumerate(hdr):
    if h in want:
        print(f"{h} [{units[i]}] = {r[i]}")
for i, h in enumerate(hdr):
    if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio") and float(r[i] or 0) > 0.05:
        print(f"  {h.split('stalled_')[1].split('_per_issue')[0]:22s} {float(r[i]):.2f} warps/issue")

src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = [i for i, x in enumerate(rows) if x and x[0] == "Address"][0]
h2 = rows[hi]
ix = {h: i for i, h in enumerate(h2)}
data = rows[hi + 1:]


def f(x, k):
    try:
        return float(x[ix[k]])
    except Exception:
        return 0.0


tot_s = sum(f(x, "# Samples") for x in data)
tot_i = sum(f(x, "Instructions Executed") for x in data)
mx = max(f(x, "Instructions Executed") for x in data)
hot = [x for x in data if f(x, "Instructions Executed") > 0.5 * mx]
stalls = [h for h in h2 if h.startswith("stall_") and "Not Issued" not in h]
print(f"source page: {len(data)} SASS instrs, {int(tot_i)} warp-instr executed, {int(tot_s)} samples")
print(f"hot loop: {len(hot)} instrs, {sum(f(x,'Instructions Executed') for x in hot)/tot_i:.1%} of executed instrs, "
      f"{sum(f(x,'# Samples') for x in hot)/tot_s:.1%} of samples")
ops = collections.Counter()
for x in hot:
    t = x[ix["Source"]].split()
    ops[(t[1] if t[0].startswith("@") else t[0]).split(".")[0]] += 1
print("hot loop mix:", dict(ops.most_common()))
for name, grp in (("hot", hot), ("rest", [x for x in data if x not in hot])):
    agg = {s: sum(f(x, s) for x in grp) for s in stalls}
    print(name, {k[6:]: int(v) for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]})
rest = sorted([x for x in data if x not in hot], key=lambda x: -f(x, "# Samples"))
print("top non-hot-loop sample sites:")
for x in rest[:14]:
    print(f"  {int(f(x,'# Samples')):5d} samples  {int(f(x,'Instructions Executed')):8d} exec  {x[ix['Source']][:64]}")
