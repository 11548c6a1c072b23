#!/bin/bash
mkdir -p gpurun_out
export VMM_TEST_2GPU_TIMEOUT_S=150
timeout 420 python -m pytest tests/test_distributed.py -m gpu -q -rA > gpurun_out/r2_pytest_2gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_2gpu.log
tail -5 gpurun_out/r2_pytest_2gpu.log
bash tools/gpu_run_nbench.sh 2
