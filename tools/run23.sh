#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log | cut -c1-200
show() { python -c "
import json,sys
d=json.loads(open('$1').read().strip().splitlines()[-1])
print('$1', round(d['ms_per_step'],2), round(d['value']), d.get('kernel_ms'))"; }
run() { name=$1; wl=$2; shift 2
  env "$@" timeout 900 python bench.py --workload $wl --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/$name.json 2> gpurun_out/$name.err
  show gpurun_out/$name.json; tail -n 2 gpurun_out/$name.err; }
for sm in 160 128 96; do
run s2d_$sm lap2d_1024 SPCHOL_SMALL_MAX_M=$sm
run s64_$sm lap3d_64 SPCHOL_SMALL_MAX_M=$sm
done
run s128_160 lap3d_128 SPCHOL_SMALL_MAX_M=160
run s128_96 lap3d_128 SPCHOL_SMALL_MAX_M=96
run sbat batched_1138 X=1
