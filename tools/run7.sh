#!/bin/bash
mkdir -p gpurun_out
show() { python -c "
import json,sys
d=json.loads(open('$1').read().strip().splitlines()[-1])
print('$1', round(d['ms_per_step'],2), round(d['value']), d['kernel_ms'], d['roofline']['by_K_class'])"; }
run() { # name, workload, env...
  name=$1; wl=$2; shift 2
  env "$@" timeout 900 python bench.py --workload $wl --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/$name.json 2> gpurun_out/$name.err
  show gpurun_out/$name.json
}
run b128_default lap3d_128 X=1
run b128_la0 lap3d_128 SPCHOL_LOOKAHEAD=0
run b128_wide3000 lap3d_128 SPCHOL_WIDE_W=3000
run b128_wide1 lap3d_128 SPCHOL_WIDE_W=1
run b64_default lap3d_64 X=1
run b64_la0 lap3d_64 SPCHOL_LOOKAHEAD=0
run b64_wide3000 lap3d_64 SPCHOL_WIDE_W=3000
run b2d_default lap2d_1024 X=1
for f in gpurun_out/b*.err; do tail -n 2 $f; done
