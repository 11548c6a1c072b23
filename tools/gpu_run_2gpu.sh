#!/bin/bash
# two-GPU NCCL parity tests (sharded persistent PCG with in-kernel exchanges; then the multi-launch loop)
mkdir -p gpurun_out
export VMM_TEST_2GPU_TIMEOUT_S=150
timeout 420 python -m pytest tests/test_distributed.py -m gpu -q -rA > gpurun_out/r2_pytest_2gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_2gpu.log
tail -5 gpurun_out/r2_pytest_2gpu.log
VMM_BA_PCG_LEGACY=1 timeout 420 python -m pytest tests/test_distributed.py -m gpu -q -rA > gpurun_out/r2_pytest_2gpu_legacy.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_2gpu_legacy.log
tail -4 gpurun_out/r2_pytest_2gpu_legacy.log
