#!/bin/bash
# one-GPU evidence run: the whole -m gpu suite, the bench line, the reference arm, the BASELINE configurations, covariance timing
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -rA --durations=10 --deselect tests/test_distributed.py > gpurun_out/r2_pytest_gpu_final.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_gpu_final.log
tail -4 gpurun_out/r2_pytest_gpu_final.log
timeout 600 python bench.py > gpurun_out/r2_bench_n1_final.json 2> gpurun_out/r2_bench_n1_final.err; echo "bench rc=$?"
timeout 400 python bench.py --impl reference --steps 2 --warmup 0 > gpurun_out/r2_bench_reference_final.json 2> gpurun_out/r2_bench_reference_final.err; echo "reference rc=$?"
timeout 300 python tools/cov_time.py c3 > gpurun_out/r2_cov_time.jsonl 2> gpurun_out/r2_cov_time.err; cat gpurun_out/r2_cov_time.jsonl
for c in c5 c4 c2; do timeout 400 python tools/run_configs.py $c > gpurun_out/r2_cfg_$c.jsonl 2> gpurun_out/r2_cfg_$c.err; done
cut -c1-260 gpurun_out/r2_cfg_c5.jsonl gpurun_out/r2_cfg_c4.jsonl
python -c "
import json
d=json.loads(open('gpurun_out/r2_bench_n1_final.json').read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['cpu_baseline'], d['solve'].get('cost_check'))
d=json.loads(open('gpurun_out/r2_bench_reference_final.json').read().strip().splitlines()[-1]); print(d.get('value'), d.get('cpu_baseline'))
"
