#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log; tail -15 gpurun_out/pytest_gpu.log | cut -c1-200
timeout 300 tools/gemm_bench v1 > gpurun_out/gb_swz.jsonl 2>&1; echo rc=$?; cut -c1-140 gpurun_out/gb_swz.jsonl
show() { python -c "
import json,sys
d=json.loads(open('$1').read().strip().splitlines()[-1])
print('$1', round(d['ms_per_step'],2), round(d['value']), d['kernel_ms'], d['roofline']['by_K_class'], d.get('solve'))"; }
for sw in 1 0; do
SPCHOL_GEMM_SWIZZLE=$sw timeout 900 python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/b128_swz$sw.json 2> gpurun_out/b128_swz$sw.err
show gpurun_out/b128_swz$sw.json; tail -n 3 gpurun_out/b128_swz$sw.err
done
SPCHOL_GEMM_SWIZZLE=1 timeout 900 python bench.py --workload lap3d_64 --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/b64_swz1.json 2> gpurun_out/b64_swz1.err
show gpurun_out/b64_swz1.json
