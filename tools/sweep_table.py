"""gpurun_out/<tag>_sweep.jsonl (tools/sweep.py) -> profiles/<tag>_sweep.md + .jsonl     usage: sweep_table.py [src.jsonl] [tag] [title suffix]"""
import json, shutil, sys
src = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/r01_sweep.jsonl"
tag = sys.argv[2] if len(sys.argv) > 2 else "r01_sweep"
suffix = sys.argv[3] if len(sys.argv) > 3 else "round 1"
rows = [json.loads(l) for l in open(src) if l.startswith("{")]
bad = [r["case"] for r in rows[1:] if r.get("filter_equals_exact") is False]
assert not bad, bad
out = ["# configs[2], configs[3] (one rank's share) and the configs[4] sweep on one B200 (tools/sweep.py, " + suffix + ")", "",
       f"FP32 peak measured in the same process: {rows[0]['fp32_peak_tflops_measured']:.1f} TFLOP/s ({rows[0]['gpu']}). "
       "`frac` = K3 tests/s x 12 FLOP / that peak.",
       "`filter = exact`: counts of the guard-banded kernel equal the exact-only kernel's (checked where marked; hn = 2048 "
       "cases are timing only). Stage times: medians of CUDA-event timings inside the C ABI. For hn > 256 `count` is `k_count_tc` alone: "
       "its constants pre-pass rides in K2's launch (`hypotheses` stage, +9 us at configs[1]), so `frac` is ~3 % higher than with the pre-pass counted.", "",
       "| case | b | vn | hn | tn total | images/s | compact µs | count µs | finalize µs | K3 T tests/s | frac | in-band | filter = exact |",
       "|---|---|---|---|---|---|---|---|---|---|---|---|---|"]
for r in rows[1:]:
    st = r["stage_ms"]
    out.append(f"| {r['case']} | {r['b']} | {r['vn']} | {r['hn']} | {r.get('tn_total', '')} | {r['images_per_s']:.0f} | "
               f"{st['compact'] * 1e3:.0f} | {st['count'] * 1e3:.0f} | {st['finalize'] * 1e3:.0f} | {r['k3_tests_per_s'] / 1e12:.2f} | "
               f"{r['k3_frac_of_fp32_peak']:.3f} | {r.get('amb_frac', float('nan')):.1e} | {r.get('filter_equals_exact')} |")
open(f"profiles/{tag}.md", "w").write("\n".join(out) + "\n")
shutil.copy(src, f"profiles/{tag}.jsonl")
print("\n".join(out[5:12]))
