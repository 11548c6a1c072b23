#!/bin/bash
mkdir -p gpurun_out
timeout 900 python tools/dist_bisect.py c3/4 > gpurun_out/r2_dist_bisect_c3_4.jsonl 2> gpurun_out/r2_dist_bisect_c3_4.err
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-c2-pair --no-cost-check > gpurun_out/r2_bench6.json 2> gpurun_out/r2_bench6.err
cat gpurun_out/r2_dist_bisect_c3_4.jsonl
