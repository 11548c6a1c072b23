#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -x -q -m gpu --deselect tests/test_distributed.py -rA --durations=8 > gpurun_out/r2_pytest_gpu_final3.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_pytest_gpu_final3.log
timeout 400 python tools/run_configs.py c4 > gpurun_out/r2_cfg_c4_b.jsonl 2> gpurun_out/r2_cfg_c4_b.err; cut -c1-330 gpurun_out/r2_cfg_c4_b.jsonl
timeout 600 python bench.py > gpurun_out/r2_bench_n1_final3.json 2> gpurun_out/r2_bench_n1_final3.err; echo "bench rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/r2_bench_n1_final3.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['gpu_launches'], d['solve']['cost_check'])"
