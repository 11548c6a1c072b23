#!/bin/bash
# round-2 two-GPU run: NCCL parity tests (sharded persistent PCG with in-kernel exchanges), then C3 at 2 ranks both ways
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2_2gpu_topo.txt 2>&1
timeout 900 python -m pytest tests/test_distributed.py -m gpu -q -rA -s > gpurun_out/r2_pytest_2gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_2gpu.log
VMM_BA_PCG_LEGACY=1 timeout 900 python -m pytest tests/test_distributed.py -m gpu -q -rA -s > gpurun_out/r2_pytest_2gpu_legacy.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_2gpu_legacy.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err
VMM_BA_PCG_LEGACY=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e > gpurun_out/r2_bench_n2_legacy.json 2> gpurun_out/r2_bench_n2_legacy.err
tail -4 gpurun_out/r2_pytest_2gpu.log; tail -c 600 gpurun_out/r2_bench_n2.err
