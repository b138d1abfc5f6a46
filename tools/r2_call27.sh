#!/bin/bash
# update kernel with explicit ld.shared fragment loads and an unpredicated main loop: microbench (bitwise vs the register-staged kernel), parity, step times
mkdir -p gpurun_out
timeout 300 ./tools/gemm_bench 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    if d['variant'] != 'r2_tma': print(d['shape'], d['variant'], d['tflops'], d['maxdiff_vs_first'])
"
timeout 900 python -m pytest tests/test_chol_gpu.py -q -m gpu -x -k "laplacians or random or golden or fullsize or ill or large" 2>&1 | tail -2
run() { wl=$1; name=$2; shift 2; env "$@" timeout 600 python bench.py --workload $wl --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-also > gpurun_out/r2u_${wl}_$name.json 2> gpurun_out/r2u_err.txt || tail -3 gpurun_out/r2u_err.txt; python - <<PY
import json
d=json.load(open('gpurun_out/r2u_${wl}_$name.json')); print('$wl', '$name', round(d['ms_per_step'],3), 'frac', d['roofline']['frac'], {k:v['tflops'] for k,v in d['roofline'].get('by_K_class',{}).items()}, 'resid', d['residual']['rel_fro_estimate'], {k:v for k,v in d['kernel_ms'].items() if isinstance(v,(int,float)) and v>0.05})
PY
}
run lap3d_128 lds X=1
run lap3d_64 lds X=1
run lap2d_1024 lds X=1
