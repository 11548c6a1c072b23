#!/bin/bash
# C3 bench at N ranks, persistent sharded PCG only
N=$1
mkdir -p gpurun_out
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2_bench_n${N}_b.json 2> gpurun_out/r2_bench_n${N}_b.err
echo "rc=$?"; tail -c 200 gpurun_out/r2_bench_n${N}_b.err
python -c "
import json
d=json.loads(open('gpurun_out/r2_bench_n${N}_b.json').read().strip().splitlines()[-1]); print($N, d['value'], d['ms_per_step'], d['e2e']['value'])
k=d['kernels']['k_pcg_persistent_dist']; print(k['ms_per_pcg_iteration'], {a:round(b,1) for a,b in k['phase_us_per_iteration_cta0'].items() if b})"
