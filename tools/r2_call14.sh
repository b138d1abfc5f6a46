#!/bin/bash
mkdir -p gpurun_out
timeout 300 ./tools/gemm_bench > gpurun_out/r2m_gemm_microbench_red.jsonl 2> /dev/null; echo "microbench rc=$?"; python - <<'PY'
import json
rows=[json.loads(l) for l in open('gpurun_out/r2m_gemm_microbench_red.jsonl')]
by={}
for r in rows: by.setdefault(r['shape'],{})[r['variant']]=(r['tflops'],r['maxdiff_vs_first'])
for k,v in by.items(): print(k, v)
PY
timeout 900 python -m pytest tests/test_chol_gpu.py -q -m gpu > gpurun_out/r2m_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/r2m_pytest_gpu.log
for SW in 3 7; do
  SPCHOL_GEMM_SWIZZLE=$SW timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-also > gpurun_out/r2m_lap3d_128_swz$SW.json 2> gpurun_out/r2m_128_swz$SW.err; echo "128 swz=$SW rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2m_lap3d_128_swz$SW.json | head -1) $(grep -o '"by_K_class": {.*}}, "whole' gpurun_out/r2m_lap3d_128_swz$SW.json) $(grep -o '"seconds_host_to_host": [0-9.]*' gpurun_out/r2m_lap3d_128_swz$SW.json)"
done
SPCHOL_GEMM_SWIZZLE=7 timeout 300 python bench.py --workload lap3d_64 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r2m_lap3d_64_swz7.json 2> /dev/null; echo "64 swz7 $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2m_lap3d_64_swz7.json | head -1)"
SPCHOL_GEMM_SWIZZLE=7 timeout 300 python bench.py --workload lap2d_1024 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r2m_lap2d_swz7.json 2> /dev/null; echo "2d swz7 $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2m_lap2d_swz7.json | head -1) $(grep -o '"seconds_host_to_host": [0-9.]*' gpurun_out/r2m_lap2d_swz7.json)"
