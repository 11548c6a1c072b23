#!/bin/bash
# usage: gpu_run_ngpu.sh N  — C3 bench at N ranks: persistent sharded PCG, then the multi-launch loop
N=$1
mkdir -p gpurun_out
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2_bench_n${N}.json 2> gpurun_out/r2_bench_n${N}.err
echo "persistent rc=$?"; tail -c 300 gpurun_out/r2_bench_n${N}.err
VMM_BA_PCG_LEGACY=1 timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 --no-e2e > gpurun_out/r2_bench_n${N}_legacy.json 2> gpurun_out/r2_bench_n${N}_legacy.err
echo "legacy rc=$?"
python - <<PY
import json
for f in ("gpurun_out/r2_bench_n${N}.json", "gpurun_out/r2_bench_n${N}_legacy.json"):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1]); print(f, d["value"], d["ms_per_step"], d["e2e"])
    except Exception as ex: print(f, "unreadable", ex)
PY
