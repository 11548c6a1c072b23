#!/bin/bash
# round-2 GPU run 3: warp-autonomous PCG phases, look-ahead sweep inversion, cluster retry, tolerance schedule sweep
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -rA -s --durations=10 > gpurun_out/r2_pytest_gpu_3.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_gpu_3.log
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-c2-pair > gpurun_out/r2_bench3.json 2> gpurun_out/r2_bench3.err
VMM_BA_PCG_LEGACY=1 timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-cost-check > gpurun_out/r2_bench3_legacy.json 2> gpurun_out/r2_bench3_legacy.err
for sc in c3/16 c3/8 c3/4 c3 c2; do f=$(echo $sc | tr / _); timeout 300 python tools/diag_convergence.py $sc huber 1.0 > gpurun_out/r2_diag3_$f.jsonl 2> gpurun_out/r2_diag3_$f.err; done
tail -5 gpurun_out/r2_pytest_gpu_3.log
