"""Print the headline and the per-workload table of a bench.py JSON line (development aid)."""
import json
import sys

d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
r = d["roofline"]
print(f"{d['config']['workload']}\n  value {d['value']:.4g} {d['unit']}  {d['ms_per_step']:.3f} ms/step  kernel {r['kernel_ms']:.3f} ms  "
      f"frac {r['frac']:.3f} ({r['bound']})  e2e {d['e2e']['value']:.4g} ({d['e2e']['ms_per_step']:.3f} ms, serial {d['e2e']['serial_ms_per_step']:.3f})  "
      f"clocks {d['clocks']['sm_mhz']} {d['clocks']['reasons']}")
if "cpu_baseline" in d:
    print(f"  cpu_baseline {d['cpu_baseline']['value']:.4g} on {d['cpu_baseline']['cores']} cores")
for k, v in d.get("workloads", {}).items():
    if "error" in v:
        print(f"  {k:12s} ERROR {v['error']}")
        continue
    rr = v["roofline"]
    print(f"  {k:12s} {v['value']:.4g} traj-steps/s  kernel {v['kernel_ms']:9.3f} ms x{v['steps']:3d}  frac {rr['frac']:.3f} ({rr['bound']}, survey-flop {rr['frac_survey_flop']:.3f})  "
          f"clk {v['clocks']['sm_mhz']} {v['clocks']['reasons']}")
for k, v in d.get("training", {}).items():
    if not isinstance(v, dict):
        print(f"  training ERROR {v}")
        break
    print(f"  train {k:12s} {v['value']:.4g} {v['unit']}  {v['n_models']} models x {v['steps_per_epoch']} steps in {v['epoch_ms']:.2f} ms ({v['us_per_step']:.1f} us per step)")
