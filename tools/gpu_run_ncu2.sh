#!/bin/bash
# one --set full capture of the persistent PCG kernel (launch of the second LM iteration of a C3 solve: ~47 PCG iterations)
mkdir -p gpurun_out
CMD="python tools/solve_trace.py c3 3"
$CMD > gpurun_out/r2_ncu2_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_pcg_persistent -s 3 -c 1 -o gpurun_out/r02_prof_pcg -f $CMD > gpurun_out/r2_ncu2_run.log 2>&1
echo "rc=$?"; ls -la gpurun_out/r02_prof_pcg.ncu-rep; tail -3 gpurun_out/r2_ncu2_run.log | cut -c1-200; grep "pcg" gpurun_out/r2_ncu2_plain.log | head -5
