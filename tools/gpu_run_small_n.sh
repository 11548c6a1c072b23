#!/bin/bash
# robustness: small workloads on several ranks (few cameras per rank, ranks without boundary markers, tiny or absent coarse spaces)
N=$1
mkdir -p gpurun_out
for w in c2 c1 c3/40; do
  f=gpurun_out/r2_small_n${N}_$(echo $w | tr / _)
  timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --workload $w --steps 6 --warmup 3 --no-cost-check --no-cpu-baseline --no-c2-pair > $f.json 2> $f.err
  echo "$w rc=$?"; tail -c 200 $f.json | cut -c1-200; echo; grep -i "error\|fail" $f.err | tail -3
  python -c "
import json
try:
    d=json.loads(open('$f.json').read().strip().splitlines()[-1]); print('$w', d['value'], d['solve']['first_solve']['final_cost'], d['solve']['first_solve']['pcg_iterations'], list(d['kernels'])[:2])
except Exception as e: print('unreadable', e)"
done
