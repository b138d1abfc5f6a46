#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log; tail -6 gpurun_out/pytest_gpu.log | cut -c1-200
show() { python -c "
import json,sys
d=json.loads(open('$1').read().strip().splitlines()[-1])
print('$1', round(d['ms_per_step'],2), round(d['value']), d['kernel_ms'], d['residual']['rel_fro_estimate'])"; }
run() { name=$1; wl=$2; shift 2
  env "$@" timeout 900 python bench.py --workload $wl --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/$name.json 2> gpurun_out/$name.err
  show gpurun_out/$name.json; tail -n 2 gpurun_out/$name.err; }
run t1d1 lap3d_128 SPCHOL_TAIL=1 SPCHOL_DEFER_CB=1
run t0d1 lap3d_128 SPCHOL_TAIL=0 SPCHOL_DEFER_CB=1
run t0d0 lap3d_128 SPCHOL_TAIL=0 SPCHOL_DEFER_CB=0
run t1d0 lap3d_128 SPCHOL_TAIL=1 SPCHOL_DEFER_CB=0
run b64_t1d1 lap3d_64 SPCHOL_TAIL=1 SPCHOL_DEFER_CB=1
run b64_t0d1 lap3d_64 SPCHOL_TAIL=0 SPCHOL_DEFER_CB=1
run b2d_t1d1 lap2d_1024 SPCHOL_TAIL=1 SPCHOL_DEFER_CB=1
run b2d_t0d0 lap2d_1024 SPCHOL_TAIL=0 SPCHOL_DEFER_CB=0
