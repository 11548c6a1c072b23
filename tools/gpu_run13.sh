#!/bin/bash
mkdir -p gpurun_out
SP_SWEEP_NO_SOLVE=1 timeout 600 python tools/sp_sweep.py c3 > gpurun_out/r2_sp_timing.jsonl 2> gpurun_out/r2_sp_timing.err
cat gpurun_out/r2_sp_timing.jsonl
