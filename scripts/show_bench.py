"""Developer helper: one-line digest of a bench.py JSON line.   python scripts/show_bench.py file.json"""
import json
import sys

d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
r = d.get("roofline", {})
print("encode", d.get("value"), "frames/s", d.get("ms_per_step"), "ms | e2e", d.get("e2e", {}).get("value"),
      "| gemm frac", r.get("frac"), "whole", r.get("whole_step_frac"), r.get("family_ms_per_step"))
s = d.get("search")
if s:
    print("search", s["value"], "q/s", s["ms_per_step"], "ms | e2e", s["e2e"]["value"], "| frac", s["roofline"]["frac"],
          [(h["Q"], h["frac_of_hbm_peak"]) for h in s.get("hbm_regime", [])])
v = d.get("vit_l14")
if v:
    print("vit_l14", v["value"], "frames/s", v["ms_per_step"], "ms | whole-step frac", v["whole_step_frac"])
c = d.get("scenecut")
if c:
    print("scenecut", c["value"], "frames/s", c["ms_per_step"], "ms | hbm frac", c["roofline"]["frac"])
dt = r.get("device_timer")
if dt:
    print("gemm device timer:", dt["avg_launch_ms"], "ms/launch, frac", dt["frac"])
