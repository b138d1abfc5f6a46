#!/bin/bash
# ncu evidence on 1 GPU (64^3 Laplacian): launch list with device times, then one --set full capture of the DMMA kernel
mkdir -p gpurun_out
CMD="python bench.py --workload lap3d_64 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches_64.csv $CMD > gpurun_out/ncu1.log 2>&1
echo "launch list rc=$?"
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_gemm_dmma -s 300 -c 3 -f -o gpurun_out/prof_gemm $CMD > gpurun_out/ncu2.log 2>&1
echo "set full rc=$?"
ls -la gpurun_out/*.ncu-rep gpurun_out/launches_64.csv; tail -3 gpurun_out/ncu2.log
