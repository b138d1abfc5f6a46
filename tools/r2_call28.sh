#!/bin/bash
# update kernel: per-stage "free" mbarriers instead of a CTA barrier per K-slab (swizzle bit 8): microbench + step times
mkdir -p gpurun_out
timeout 300 ./tools/gemm_bench 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    if d['variant'].startswith('r2_tma_red'): print(d['shape'], d['variant'], d['tflops'], d['maxdiff_vs_first'])
"
run() { wl=$1; name=$2; shift 2; env "$@" timeout 600 python bench.py --workload $wl --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-also > gpurun_out/r2v_${wl}_$name.json 2> gpurun_out/r2v_err.txt || tail -3 gpurun_out/r2v_err.txt; python - <<PY
import json
d=json.load(open('gpurun_out/r2v_${wl}_$name.json')); print('$wl', '$name', round(d['ms_per_step'],3), 'frac', d['roofline']['frac'], {k:v['tflops'] for k,v in d['roofline'].get('by_K_class',{}).items()}, 'resid', d['residual']['rel_fro_estimate'])
PY
}
run lap3d_128 swz7 SPCHOL_GEMM_SWIZZLE=7
run lap3d_128 swz15 SPCHOL_GEMM_SWIZZLE=15
run lap3d_64 swz15 SPCHOL_GEMM_SWIZZLE=15
run lap2d_1024 swz15 SPCHOL_GEMM_SWIZZLE=15
