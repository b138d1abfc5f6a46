#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline"
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -s 4570 -c 1530 --csv --log-file gpurun_out/launches_128.csv $CMD > gpurun_out/ncu_l128.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/launches_128.csv
