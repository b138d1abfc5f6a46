#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log; tail -6 gpurun_out/pytest_gpu.log | cut -c1-200
timeout 1200 python bench.py --no-cpu-baseline > gpurun_out/bench_e2e.json 2> gpurun_out/bench_e2e.err; echo "bench rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/bench_e2e.json').read().strip().splitlines()[-1]); print(round(d['ms_per_step'],1), d['e2e'], d['residual']['rel_fro_estimate'], d['solve'])"
tail -n 3 gpurun_out/bench_e2e.err
