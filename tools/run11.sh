#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
CMD="python bench.py --workload batched_1138 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/batched_plain.json 2> gpurun_out/batched_plain.err; echo rc=$?; cut -c1-300 gpurun_out/batched_plain.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 486 -c 170 --csv --log-file gpurun_out/launches_batched.csv $CMD > gpurun_out/ncu_batched.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_batched.log | cut -c1-200; wc -l gpurun_out/launches_batched.csv
timeout 900 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/b128_resid.json 2> gpurun_out/b128_resid.err; echo rc=$?
python -c "
import json
d=json.loads(open('gpurun_out/b128_resid.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['residual'])"
tail -n 3 gpurun_out/b128_resid.err
