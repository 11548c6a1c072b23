#!/bin/bash
# round-2 two-GPU run: NCCL parity tests (sharded persistent PCG with in-kernel exchanges), then C3 at 2 ranks both ways
mkdir -p gpurun_out
export VMM_TEST_2GPU_TIMEOUT_S=150
timeout 420 python -m pytest tests/test_distributed.py -m gpu -q -rA -s > gpurun_out/r2_pytest_2gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_2gpu.log
tail -5 gpurun_out/r2_pytest_2gpu.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err
tail -c 400 gpurun_out/r2_bench_n2.err; tail -c 300 gpurun_out/r2_bench_n2.json
