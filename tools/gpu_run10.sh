#!/bin/bash
mkdir -p gpurun_out
DIST_BISECT_ONLY="plain kernel,split every 2,split every 1000,split every 2, legacy loop" timeout 600 python tools/dist_bisect.py c3/4 > gpurun_out/r2_dist_fix.jsonl 2> gpurun_out/r2_dist_fix.err
cat gpurun_out/r2_dist_fix.jsonl
if grep -q "FAIL\|TIMEOUT" gpurun_out/r2_dist_fix.jsonl; then
  timeout 120 python tools/dist_trace.py dbg_trace gpurun_out/r2_trace_fix.npy > gpurun_out/r2_trace_fix.log 2>&1; cat gpurun_out/r2_trace_fix.log
else
  timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-c2-pair --no-cost-check > gpurun_out/r2_bench10.json 2> gpurun_out/r2_bench10.err; tail -c 600 gpurun_out/r2_bench10.json
fi
