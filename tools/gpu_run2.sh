#!/bin/bash
# round-2 GPU run 2: retuned persistent PCG (phase clock), new defaults, block covariance, coarse inverse diagnostics
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -rA -s --durations=15 > gpurun_out/r2_pytest_gpu_2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_gpu_2.log
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench2.json 2> gpurun_out/r2_bench2.err
VMM_BA_PCG_LEGACY=1 timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-cost-check > gpurun_out/r2_bench2_legacy.json 2> gpurun_out/r2_bench2_legacy.err
timeout 300 python tools/diag_convergence.py c3/8 huber 1.0 > gpurun_out/r2_diag2_c3_8.jsonl 2> gpurun_out/r2_diag2_c3_8.err
timeout 300 python tools/diag_convergence.py c3/16 huber 1.0 > gpurun_out/r2_diag2_c3_16.jsonl 2> gpurun_out/r2_diag2_c3_16.err
timeout 300 python tools/diag_convergence.py c3/4 huber 1.0 > gpurun_out/r2_diag2_c3_4.jsonl 2> gpurun_out/r2_diag2_c3_4.err
timeout 300 python tools/diag_convergence.py c2 huber 1.0 > gpurun_out/r2_diag2_c2.jsonl 2> gpurun_out/r2_diag2_c2.err
timeout 300 python tools/diag_convergence.py c3 huber 1.0 > gpurun_out/r2_diag2_c3.jsonl 2> gpurun_out/r2_diag2_c3.err
tail -5 gpurun_out/r2_pytest_gpu_2.log
