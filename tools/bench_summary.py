"""Collect bench.py JSON lines (gpurun_out/bench_*.json or a profiles/ directory) into one markdown table.
python tools/bench_summary.py DIR > DIR/BENCH_SUMMARY.md"""
import glob, json, os, sys
d = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out"
rows = []
for f in sorted(glob.glob(os.path.join(d, "bench_*.json"))):
    try:
        line = [l for l in open(f) if l.startswith("{")][-1]
        j = json.loads(line)
    except Exception:
        continue
    if j.get("impl") == "reference":
        rows.append((os.path.basename(f), "reference (CPU port)", j["value"], j["ms_per_step"], None, None, None, j.get("cpu_baseline", {}).get("cores")))
        continue
    r = j["roofline"]
    rows.append((os.path.basename(f), j["dtype"].split(" ")[0], j["value"], j["ms_per_step"], r["achieved"], f'{r["bound"]} {r["frac"]:.2f}',
                 j["e2e"]["value"], j["clocks"].get("sm_mhz")))
print("| file | arithmetic | traj-steps/s (device) | ms/step | TFLOP/s (algorithmic) | roofline frac | traj-steps/s (e2e) | SM MHz / cores |")
print("|---|---|---|---|---|---|---|---|")
for a, b, c, dd, e, f, g, h in rows:
    print(f"| {a} | {b} | {c:.4g} | {dd:.4g} | {'' if e is None else f'{e:.4g}'} | {f or ''} | {'' if g is None else f'{g:.4g}'} | {h} |")
