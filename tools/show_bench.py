"""Developer helper: the numbers of a bench.py JSON line worth reading at a glance."""
import json
import sys

d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("value %.0f %s  ms/step %.4f  stage_ms %s" % (d["value"], d["unit"], d["ms_per_step"], d.get("stage_ms")))
r = d.get("roofline", {})
print("roofline frac %.3f achieved %.0f GB/s (dense kernel %.4f ms)  e2e %.0f  clocks %s" % (
    r.get("frac", 0), r.get("achieved", 0), r.get("avg_launch_ms", 0), d["e2e"]["value"], d.get("clocks")))
print("config", {k: d["config"][k] for k in ("candidates_per_image", "kept_per_image") if k in d["config"]},
      "cpu_baseline", d.get("cpu_baseline"))
