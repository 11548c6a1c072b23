#!/bin/bash
# eight ranks: C3 bench (persistent sharded PCG, e2e included), the multi-launch loop beside it, then C5 (R+J sweep) and C4 (Huber)
N=${1:-8}
mkdir -p gpurun_out
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2_bench_n${N}.json 2> gpurun_out/r2_bench_n${N}.err
echo "persistent rc=$?"; tail -c 300 gpurun_out/r2_bench_n${N}.err
VMM_BA_PCG_LEGACY=1 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 --no-e2e > gpurun_out/r2_bench_n${N}_legacy.json 2> gpurun_out/r2_bench_n${N}_legacy.err
echo "legacy rc=$?"
C4_HUBER_ONLY=1 timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 tools/run_configs_dist.py c5c4 64000000 > gpurun_out/r2_c5c4_n$N.jsonl 2> gpurun_out/r2_c5c4_n$N.err
echo "c5c4 rc=$?"; cat gpurun_out/r2_c5c4_n$N.jsonl | cut -c1-330; tail -3 gpurun_out/r2_c5c4_n$N.err
python - <<PY
import json
for f in ("gpurun_out/r2_bench_n${N}.json", "gpurun_out/r2_bench_n${N}_legacy.json"):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1]); print(f, d["value"], d["ms_per_step"], (d.get("e2e") or {}).get("value"))
    except Exception as ex: print(f, "unreadable", ex)
PY
