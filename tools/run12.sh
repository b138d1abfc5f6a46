#!/bin/bash
mkdir -p gpurun_out
echo skip-pytest

for ch in 1 0; do
SPCHOL_BATCH_CHAIN=$ch timeout 600 python bench.py --workload batched_1138 --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/batched_ch$ch.json 2> gpurun_out/batched_ch$ch.err; echo rc=$?
python -c "
import json
d=json.loads(open('gpurun_out/batched_ch$ch.json').read().strip().splitlines()[-1]); print('chain=$ch', d['ms_per_step'], d['value'], d['roofline']['frac'], d['gpu_launches'])"
tail -n 2 gpurun_out/batched_ch$ch.err
done
CMD="python bench.py --workload batched_1138 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/launches_batched2.csv $CMD > gpurun_out/ncu_batched2.log 2>&1
echo "ncu rc=$?"
