#!/bin/bash
# round-2 GPU run 5: split-marker path under compute-sanitizer first, tests, bench (+ spatial order A/B), SP sweep, diagnostics
mkdir -p gpurun_out
timeout 300 compute-sanitizer --tool memcheck --launch-timeout 60 python tools/solve_split.py > gpurun_out/r2_sanitizer_split.log 2>&1; echo "rc=$?" >> gpurun_out/r2_sanitizer_split.log
timeout 1500 python -m pytest tests -m gpu -q -rA -s --durations=10 > gpurun_out/r2_pytest_gpu_5.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_gpu_5.log
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-c2-pair > gpurun_out/r2_bench5.json 2> gpurun_out/r2_bench5.err
VMM_BA_SPATIAL_ORDER=1 timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-c2-pair --no-cost-check > gpurun_out/r2_bench5_spatial.json 2> gpurun_out/r2_bench5_spatial.err
VMM_BA_PCG_LEGACY=1 timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-cost-check > gpurun_out/r2_bench5_legacy.json 2> gpurun_out/r2_bench5_legacy.err
VMM_BA_LIBRARY=$PWD/tools/variants/libvmm_ba_invv2.so timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "coarse_inverse or covariance_matches_dense" > gpurun_out/r2_inv5_invv2.log 2>&1
timeout 900 python tools/sp_sweep.py c3 > gpurun_out/r2_sp_sweep5.jsonl 2> gpurun_out/r2_sp_sweep5.err
for sc in c3/16 c3/8 c3/4 c3 c2; do f=$(echo $sc | tr / _); timeout 300 python tools/diag_convergence.py $sc huber 1.0 > gpurun_out/r2_diag5_$f.jsonl 2> gpurun_out/r2_diag5_$f.err; done
tail -5 gpurun_out/r2_pytest_gpu_5.log
