#!/bin/bash
N=$1; MAXC=$2; C4DIV=$3
mkdir -p gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 tools/run_configs_dist.py c5 $MAXC > gpurun_out/r2_c5_n$N.jsonl 2> gpurun_out/r2_c5_n$N.err
echo "c5 rc=$?"; cat gpurun_out/r2_c5_n$N.jsonl; tail -3 gpurun_out/r2_c5_n$N.err
if [ -n "$C4DIV" ]; then
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 tools/run_configs_dist.py c4 $C4DIV > gpurun_out/r2_c4_n$N.jsonl 2> gpurun_out/r2_c4_n$N.err
echo "c4 rc=$?"; cat gpurun_out/r2_c4_n$N.jsonl; tail -3 gpurun_out/r2_c4_n$N.err
fi
