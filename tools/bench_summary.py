"""Print the headline numbers and the per-kernel table of a bench.py JSON line (file may hold other lines)."""
import json, sys
for l in open(sys.argv[1]):
    l = l.strip()
    if not l.startswith("{"):
        continue
    d = json.loads(l)
    print({k: d.get(k) for k in ("value", "ms_per_step", "n_gpus", "gpu_launches")}, "e2e", (d.get("e2e") or {}).get("value"))
    print("roofline", d.get("roofline"))
    for k, v in sorted(d.get("kernels", {}).items(), key=lambda kv: -kv[1]["ms_total"])[:int(sys.argv[2]) if len(sys.argv) > 2 else 16]:
        print(f"{k:36s} n={v['launches']:5d} avg={v['ms_avg']:.4f} share={v['share']:.3f}")
    print(d.get("solve"))
