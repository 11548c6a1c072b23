#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e > gpurun_out/r2_bench_n2_b.json 2> gpurun_out/r2_bench_n2_b.err
tail -c 300 gpurun_out/r2_bench_n2_b.err; tail -c 200 gpurun_out/r2_bench_n2_b.json
