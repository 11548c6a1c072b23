#!/bin/bash
# launch list of the bench command (cold-cache, serialised: shares, not absolutes)
mkdir -p gpurun_out
CMD="python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-c2-pair --no-cost-check --no-e2e"
$CMD > gpurun_out/r2_ncu_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r02_launches.csv $CMD > gpurun_out/r2_ncu_launches_run.log 2>&1
echo "rc=$?"; wc -l gpurun_out/r02_launches.csv; tail -2 gpurun_out/r2_ncu_launches_run.log | cut -c1-300
