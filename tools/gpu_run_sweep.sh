#!/bin/bash
mkdir -p gpurun_out
timeout 900 python tools/sp_sweep.py ${1:-c3} > gpurun_out/r2_sp_sweep_final.jsonl 2> gpurun_out/r2_sp_sweep_final.err
cat gpurun_out/r2_sp_sweep_final.jsonl | cut -c1-330
