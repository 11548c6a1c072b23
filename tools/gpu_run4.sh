#!/bin/bash
# round-2 GPU run 4: tests with the checked sweep inversion, bench, inversion variants on the device, SP knob sweep
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -rA -s --durations=10 > gpurun_out/r2_pytest_gpu_4.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_gpu_4.log
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-c2-pair > gpurun_out/r2_bench4.json 2> gpurun_out/r2_bench4.err
VMM_BA_PCG_LEGACY=1 timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-cost-check > gpurun_out/r2_bench4_legacy.json 2> gpurun_out/r2_bench4_legacy.err
for v in invv2 invlook; do
  VMM_BA_LIBRARY=$PWD/tools/variants/libvmm_ba_$v.so timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "coarse_inverse or covariance_matches_dense" > gpurun_out/r2_inv_$v.log 2>&1
done
timeout 900 python tools/sp_sweep.py c3 > gpurun_out/r2_sp_sweep.jsonl 2> gpurun_out/r2_sp_sweep.err
for sc in c3/16 c3/8 c3/4 c3 c2; do f=$(echo $sc | tr / _); timeout 300 python tools/diag_convergence.py $sc huber 1.0 > gpurun_out/r2_diag4_$f.jsonl 2> gpurun_out/r2_diag4_$f.err; done
tail -5 gpurun_out/r2_pytest_gpu_4.log
