#!/bin/bash
# tier-2 subtree tasks: parity subset, then 2D / 3D timings with the tier on (default) and off, NW of the tier
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_chol_gpu.py -q -m gpu -k "laplacians or random or subtree or golden or fullsize" 2>&1 | tail -2
run() { wl=$1; name=$2; shift 2; env "$@" timeout 600 python bench.py --workload $wl --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-also > gpurun_out/r2q_${wl}_$name.json 2> gpurun_out/r2q_err.txt || tail -3 gpurun_out/r2q_err.txt; python - <<PY
import json
d=json.load(open('gpurun_out/r2q_${wl}_$name.json')); print('$wl', '$name', round(d['ms_per_step'],3), 'resid', d.get('residual'), {k:v for k,v in d['kernel_ms'].items() if isinstance(v,(int,float)) and v>0.05})
PY
}
run lap2d_1024 tier2 X=1
run lap2d_1024 off SPCHOL_TREE2_CAP=0
run lap2d_1024 nw2 SPCHOL_TREE2_NW=2
run lap2d_1024 cap96 SPCHOL_TREE2_CAP=96
run lap2d_1024 work3e6 SPCHOL_TREE2_WORK=3e6
run lap2d_1024 t1cap64 SPCHOL_TREE_CAP=64
run lap3d_64 tier2 X=1
run lap3d_64 off SPCHOL_TREE2_CAP=0
run lap3d_128 tier2 X=1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2q_launches_2d.csv python bench.py --workload lap2d_1024 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-also > /dev/null 2>&1
python tools/summarize_launches.py gpurun_out/r2q_launches_2d.csv gpurun_out/r2q_launches_2d.txt "bench.py --workload lap2d_1024 --steps 1 --warmup 3 (6 factorizations + residual + 2 solves)" | head -24
