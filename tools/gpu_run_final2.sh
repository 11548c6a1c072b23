#!/bin/bash
# final one-GPU run of the round: whole -m gpu suite, then the default bench line
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -x -q -m gpu --deselect tests/test_distributed.py -rA --durations=8 > gpurun_out/r2_pytest_gpu_final2.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_pytest_gpu_final2.log
timeout 600 python bench.py > gpurun_out/r2_bench_n1_final2.json 2> gpurun_out/r2_bench_n1_final2.err; echo "bench rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/r2_bench_n1_final2.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['gpu_launches'], d['solve']['cost_check']['default_vs_oracle_rel'])
k=d['kernels']['k_pcg_persistent']; print(k['ms_avg'], k['ms_per_pcg_iteration'], k['phase_us_per_iteration_cta0'])"
