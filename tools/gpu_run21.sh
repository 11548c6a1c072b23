#!/bin/bash
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_site_scale.py -m gpu -q -k "ring_laps" -rA > gpurun_out/r2_pytest_ringlap.log 2>&1; tail -3 gpurun_out/r2_pytest_ringlap.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2_smoke.log 2>&1; tail -2 gpurun_out/r2_smoke.log
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-c2-pair --no-cost-check --pcg-tolerance 5e-2 > gpurun_out/r2_bench_tol5e-2.json 2> gpurun_out/r2_bench_tol5e-2.err
python -c "
import json
d=json.loads(open('gpurun_out/r2_bench_tol5e-2.json').read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['solve']['first_solve'])"
