#!/bin/bash
mkdir -p gpurun_out
for v in default inl; do
  if [ $v = default ]; then unset VMM_BA_LIBRARY; else export VMM_BA_LIBRARY=$PWD/tools/variants/libvmm_ba_$v.so; fi
  timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-c2-pair --no-cost-check --no-e2e > gpurun_out/r2_bench15_$v.json 2> gpurun_out/r2_bench15_$v.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/r2_bench15_$v.json").read().strip().splitlines()[-1])
k=d["kernels"]["k_pcg_persistent"]
print("$v", d["value"], d["ms_per_step"], k["ms_per_pcg_iteration"], {a:round(b,1) for a,b in k["phase_us_per_iteration_cta0"].items()})
PY
done
