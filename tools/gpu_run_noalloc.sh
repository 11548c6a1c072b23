#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -rA > gpurun_out/r2_pytest_noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -4 gpurun_out/r2_pytest_noalloc_alone.log
timeout 1200 python -m pytest tests -x -q -m gpu --deselect tests/test_distributed.py -rA --durations=5 > gpurun_out/r2_pytest_gpu_all.log 2>&1; echo "all rc=$?"; tail -6 gpurun_out/r2_pytest_gpu_all.log
