#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --workload lap2d_1024 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_2d.csv $CMD > gpurun_out/ncu_l2d.log 2>&1
echo "rc=$?"; wc -l gpurun_out/launches_2d.csv
