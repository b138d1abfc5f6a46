#!/bin/bash
# launch list of config 3 (2D 1024^2) on the final code: the two tiers of subtree tasks next to the level program
mkdir -p gpurun_out
timeout 150 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2z_launches_lap2d_1024.csv python bench.py --workload lap2d_1024 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-also > /dev/null 2>&1
echo "rc=$?"; wc -l gpurun_out/r2z_launches_lap2d_1024.csv
