#!/bin/bash
# the last GPU minute of round 1: path groups as the default batched kernel -- every batched GPU test, one bench line,
# one --set full capture for the next round
mkdir -p gpurun_out
timeout 100 python -m pytest tests/test_chol_gpu.py -q -x -k "batched" > gpurun_out/batched_tests.log 2>&1
echo "batched tests rc=$?"; tail -n 4 gpurun_out/batched_tests.log
timeout 100 ncu --set full --clock-control none --import-source on -k regex:k_front_path -c 1 -f -o gpurun_out/prof_front_path_v2 \
  python bench.py --workload batched_1138 --nbatch 2048 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/ncu_front_path_v2.log 2>&1
echo "ncu rc=$?"
