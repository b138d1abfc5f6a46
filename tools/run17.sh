#!/bin/bash
mkdir -p gpurun_out
for nb in 148 444 888 4440; do
timeout 300 python bench.py --workload batched_1138 --nbatch $nb --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/bt_$nb.json 2> gpurun_out/bt_$nb.err
python -c "
import json
d=json.loads(open('gpurun_out/bt_$nb.json').read().strip().splitlines()[-1]); print('nbatch=$nb', round(d['ms_per_step'],3))"
done
