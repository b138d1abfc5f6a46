#!/bin/bash
mkdir -p gpurun_out
show() { python -c "
import json,sys
d=json.loads(open('$1').read().strip().splitlines()[-1])
print('$1', round(d['ms_per_step'],2), round(d['value']), d['kernel_ms'], d['roofline']['by_K_class'])"; }
for sw in 3 1 3 1; do
SPCHOL_GEMM_SWIZZLE=$sw timeout 900 python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/b128_pf$sw.json 2> gpurun_out/b128_pf$sw.err
show gpurun_out/b128_pf$sw.json; tail -n 3 gpurun_out/b128_pf$sw.err
done
