#!/bin/bash
mkdir -p gpurun_out
timeout 300 tools/gemm_bench v1 > gpurun_out/gb_v1.jsonl 2>&1; echo rc=$?
ncu --set full --clock-control none --import-source on -k regex:k_gemm_dmma -s 2 -c 1 -f -o gpurun_out/prof_gemm_micro tools/gemm_bench v1 > gpurun_out/ncu_micro.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_micro.log; ls -la gpurun_out/*.ncu-rep
