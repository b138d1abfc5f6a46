#!/bin/bash
# one --set full capture of the staged path-group kernel (2048 matrices keep the replay short)
mkdir -p gpurun_out
SPCHOL_BATCH_PATH=1 timeout 150 ncu --set full --clock-control none --import-source on -k regex:k_front_path -c 1 -f -o gpurun_out/prof_front_path \
  python bench.py --workload batched_1138 --nbatch 2048 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/ncu_front_path.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/prof_front_path.ncu-rep 2>/dev/null; tail -n 3 gpurun_out/ncu_front_path.log
