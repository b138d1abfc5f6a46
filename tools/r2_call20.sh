#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_ingest_gpu.py tests/test_chol_gpu.py -q -m gpu -k "ingest or solve or laplacians or random" 2>&1 | tail -2
run() { name=$1; shift; env "$@" timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-also > gpurun_out/r2p_128_$name.json 2> /dev/null; python - <<PY
import json
d=json.load(open('gpurun_out/r2p_128_$name.json')); print('$name', round(d['ms_per_step'],1), d['roofline']['frac'], {k:(v['ms'],v['tflops']) for k,v in d['roofline']['by_K_class'].items()}, 'solve', d['solve']['seconds_host_to_host'], 'potrf', d['kernel_ms']['potrf'], 'trsm', d['kernel_ms']['trsm'])
PY
}
run base X=1
run nb512 SPCHOL_NB=512
run wide1024 SPCHOL_NB_WIDE=1024
run widew4096 SPCHOL_WIDE_W=4000
run nb512_wide1024 SPCHOL_NB=512 SPCHOL_NB_WIDE=1024
timeout 300 python bench.py --workload lap3d_64 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r2p_64.json 2>/dev/null; python - <<'PY'
import json
d=json.load(open('gpurun_out/r2p_64.json')); print('64^3', round(d['ms_per_step'],2), 'solve', d['solve'])
PY
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2p_launches_2d.csv python bench.py --workload lap2d_1024 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > /dev/null 2>&1
python tools/summarize_launches.py gpurun_out/r2p_launches_2d.csv gpurun_out/r2p_launches_2d.txt "bench.py --workload lap2d_1024 --steps 1 --warmup 3 (6 factorizations + residual + 2 solves)" | head -24
