#!/bin/bash
# round-2 GPU run 1: site-scale parity tests, convergence diagnostics, slot-release A/B
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/r2_run1_env.txt; nproc >> gpurun_out/r2_run1_env.txt; free -g >> gpurun_out/r2_run1_env.txt; lscpu | grep "Model name" >> gpurun_out/r2_run1_env.txt
# the persistent PCG kernel first, on two small solves: if it misbehaves the rest of the run uses the round-1 loop
if ! timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "solve_parity_c1 or zero_noise or solve_parity_c2" > gpurun_out/r2_pcg_smoke.log 2>&1; then export VMM_BA_PCG_LEGACY=1; echo "LEGACY PCG" >> gpurun_out/r2_pcg_smoke.log; fi
timeout 1500 python -m pytest tests -m gpu -q -rA -s --durations=15 > gpurun_out/r2_pytest_gpu_1.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_gpu_1.log
timeout 300 python tools/diag_convergence.py c3 huber 1.0 > gpurun_out/r2_diag_c3.jsonl 2> gpurun_out/r2_diag_c3.err
timeout 300 python tools/diag_convergence.py c3/8 huber 1.0 > gpurun_out/r2_diag_c3_8.jsonl 2> gpurun_out/r2_diag_c3_8.err
timeout 600 python tools/diag_convergence.py c4 huber 1.0 > gpurun_out/r2_diag_c4.jsonl 2> gpurun_out/r2_diag_c4.err
VMM_BA_PCG_LEGACY=1 timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/r2_bench_legacy.json 2> gpurun_out/r2_bench_legacy.err
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2_bench_fence.json 2> gpurun_out/r2_bench_fence.err
for m in 0 2; do
  VMM_BA_LIBRARY=$PWD/tools/variants/libvmm_ba_rel$m.so timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/r2_bench_rel$m.json 2> gpurun_out/r2_bench_rel$m.err
done
tail -5 gpurun_out/r2_pytest_gpu_1.log
