#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log; tail -6 gpurun_out/pytest_gpu.log | cut -c1-200
show() { python -c "
import json,sys
d=json.loads(open('$1').read().strip().splitlines()[-1])
print('$1', round(d['ms_per_step'],2), round(d['value']), d.get('kernel_ms'))"; }
run() { name=$1; wl=$2; shift 2
  timeout 900 python bench.py --workload $wl --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/$name.json 2> gpurun_out/$name.err
  show gpurun_out/$name.json; tail -n 2 gpurun_out/$name.err; }
run p2d lap2d_1024
run p64 lap3d_64
run p128 lap3d_128
run pbat batched_1138
