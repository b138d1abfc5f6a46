#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "batched" > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
for ch in 1; do
SPCHOL_BATCH_CHAIN=$ch timeout 600 python bench.py --workload batched_1138 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/batched_ch$ch.json 2> gpurun_out/batched_ch$ch.err; echo rc=$?
python -c "
import json
d=json.loads(open('gpurun_out/batched_ch$ch.json').read().strip().splitlines()[-1]); print('chain=$ch', d['ms_per_step'], d['value'], d['roofline']['frac'], d['gpu_launches'], d['e2e'])"
tail -n 2 gpurun_out/batched_ch$ch.err
done
