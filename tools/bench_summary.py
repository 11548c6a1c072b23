"""Print the headline and the per-kernel table of a bench.py JSON line."""
import json, sys
d = json.load(open(sys.argv[1]))
print("value", round(d["value"], 3), d["unit"], "ms/step", round(d["ms_per_step"], 3), "e2e", d.get("e2e") and round(d["e2e"]["value"], 3))
print("roofline", d["roofline"])
print("rj", d["rj"]["roofline"]["frac"], d["rj"]["ms_per_pass"])
for k, v in list(d["kernels"].items())[:10]:
    print(f"  {k:34s} n={v['launches']:5d} avg={v['ms_avg']:.4f} ms share={v['share']:.3f} frac_hbm={v.get('frac_hbm', 0):.3f}")
print(d["solve"])
