"""print the interesting parts of a bench.py JSON line"""
import json, sys
d = json.load(open(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/bench.json"))
r = d.get("roofline", {})
print("value %.4g %s  ms/step %.3f  e2e %.4g (%.1f ms/step)  launches %s" % (d["value"], d["unit"], d["ms_per_step"],
      d["e2e"]["value"] or 0, d["e2e"].get("ms_per_step", 0), d.get("gpu_launches")))
print("roofline frac %.4f achieved %.1f kernel %s | iteration frac %.4f" % (r.get("frac", 0), r.get("achieved", 0), r.get("kernel"),
      r.get("iteration", {}).get("frac", 0)))
for k, v in sorted(r.get("kernels", {}).items(), key=lambda kv: -kv[1]["total_ms"]):
    print("  %-22s n=%3d avg %9.1f us total %8.3f ms" % (k, v["launches"], v["avg_us"], v["total_ms"]))
for k, v in d.get("others", {}).items():
    print("  other", k, {a: (round(b, 3) if isinstance(b, float) else b) for a, b in v.items() if a != "note"})
print("cpu", d.get("cpu_baseline"), "clocks", d.get("clocks"))
